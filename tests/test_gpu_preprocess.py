"""K1 parity: bt_preprocess_launch (through the C ABI) against the oracle and against the
outputs of the unmodified reference (tests/golden/rec_small_expected.npz).

Bars (BASELINE.json north_star): bit-exact for uint8 work, flips and transposes; float32 within
1e-6 relative -- here it is in fact bit-exact, which is asserted; bfloat16 within 1 ulp (asserted
0 ulp against round-to-nearest-even of the float32 oracle)."""
import itertools

import numpy as np
import pytest
import torch

from blendtorch import btt
from oracle import btr, preprocess as opre

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _run(frames, **kw):
    spec = btt.Preprocess(**kw)
    src = torch.from_numpy(np.ascontiguousarray(frames)).to(DEV)
    out = btt.preprocess_images(src, spec)
    if spec.stats:
        out, st = out
        torch.cuda.synchronize()
        return out, st
    torch.cuda.synchronize()
    return out


def _oracle(frames, flip, gamma, mean, std, layout, dtype, out_channels=None):
    out_dtype = "uint8" if dtype == torch.uint8 else "float32"
    exp = opre.preprocess_batch(list(frames), flip=flip, gamma=gamma, mean=mean, std=std,
                                chw=(layout == "NCHW"), out_dtype=out_dtype)
    if out_channels:
        exp = exp[:, :out_channels] if layout == "NCHW" else exp[..., :out_channels]
    return np.ascontiguousarray(exp)


def _assert_matches(out, exp, dtype):
    if dtype == torch.bfloat16:
        got = out.view(torch.int16).cpu().numpy().view(np.uint16)
        want = opre.to_bf16_bits(exp)
        ulp = np.abs(got.astype(np.int32) - want.astype(np.int32)).max() if got.size else 0
        assert ulp <= 1, "bf16 differs by %d ulp" % ulp   # tolerance stated by north_star
        assert ulp == 0
    else:
        got = out.cpu().numpy()
        assert got.dtype == exp.dtype and got.shape == exp.shape
        if got.dtype == np.float32:
            # bit-exact, which implies the 1e-6 relative bar
            np.testing.assert_array_equal(got.view(np.uint32), exp.view(np.uint32))
        else:
            np.testing.assert_array_equal(got, exp)


SHAPES = [(2, 12, 16, 4), (3, 7, 5, 4), (2, 9, 16, 3), (2, 8, 32, 3), (1, 5, 16, 1), (2, 6, 10, 2),
          (1, 4, 8, 4), (2, 3, 24, 4), (5, 1, 4, 4), (1, 33, 48, 3)]


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.uint8])
def test_parity_sweep(shape, dtype):
    rng = np.random.default_rng(hash(shape) % 2 ** 32)
    frames = rng.integers(0, 256, shape, dtype=np.uint8)
    C = shape[3]
    norms = [(None, None)] if dtype == torch.uint8 else [
        (None, None), (127.5, 127.5), ([0.485 * 255, 0.456 * 255, 0.406 * 255, 3.0][:C], [58.4, 57.1, 57.4, 2.0][:C])]
    for flip, gamma, (mean, std), layout in itertools.product((False, True), (None, 2.2), norms, ("NCHW", "NHWC")):
        out = _run(frames, flip=flip, gamma=gamma, mean=mean, std=std, layout=layout, dtype=dtype)
        _assert_matches(out, _oracle(frames, flip, gamma, mean, std, layout, dtype), dtype)
    if C == 4:
        out = _run(frames, gamma=2.2, mean=127.5, std=127.5, dtype=dtype if dtype != torch.uint8 else torch.float32,
                   out_channels=3)
        _assert_matches(out, _oracle(frames, False, 2.2, 127.5, 127.5, "NCHW", torch.float32, 3),
                        dtype if dtype != torch.uint8 else torch.float32)


def test_against_reference_fixture(golden_dir):
    exp = np.load(golden_dir / "rec_small_expected.npz")
    ds = btr.ReplayDataset(str(golden_dir / "rec_small"))
    frames = np.stack([ds[i]["image"] for i in range(len(ds))])
    want = np.concatenate([exp["gamma_norm_b%d_image" % k] for k in range(3)])
    got = _run(frames, gamma=2.2, mean=127.5, std=127.5).cpu().numpy()
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
    want = np.concatenate([exp["norm_b%d_image" % k] for k in range(3)])
    np.testing.assert_array_equal(_run(frames, mean=127.5, std=127.5).cpu().numpy(), want)
    want = np.concatenate([exp["gamma_b%d_image" % k] for k in range(3)])
    np.testing.assert_array_equal(_run(frames, gamma=2.2, layout="NHWC", dtype=torch.uint8).cpu().numpy(), want)
    want = np.concatenate([exp["offscreen_b%d_image" % k] for k in range(3)])
    np.testing.assert_array_equal(_run(frames, flip=True, gamma=2.2, layout="NHWC", dtype=torch.uint8).cpu().numpy(), want)
    want = np.concatenate([exp["offscreen_rgb_b%d_image" % k] for k in range(3)])
    got = _run(np.ascontiguousarray(frames[..., :3]), flip=True, gamma=2.2, layout="NHWC", dtype=torch.uint8)
    np.testing.assert_array_equal(got.cpu().numpy(), want)


def test_variants_agree_and_vector_path_is_taken():
    from blendtorch.btt import native as N
    rng = np.random.default_rng(5)
    for shape in ((4, 48, 64, 4), (2, 16, 64, 3), (3, 10, 32, 1)):
        frames = rng.integers(0, 256, shape, dtype=np.uint8)
        for dtype in (torch.float32, torch.bfloat16, torch.uint8):
            kw = dict(flip=True, gamma=2.2, dtype=dtype)
            if dtype != torch.uint8:
                kw.update(mean=127.5, std=127.5)
            a = _run(frames, variant="generic", **kw)
            assert N.lib.bt_preprocess_last_variant() == 1
            b = _run(frames, variant="vector", **kw)
            assert N.lib.bt_preprocess_last_variant() == 2
            c = _run(frames, **kw)
            assert N.lib.bt_preprocess_last_variant() == 2     # auto picks the vector kernel
            assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
            assert torch.equal(a.view(torch.uint8), c.view(torch.uint8))
    with pytest.raises(N.NativeError):
        _run(rng.integers(0, 256, (1, 4, 5, 4), dtype=np.uint8), variant="vector")


def test_odd_widths_take_the_aligned_group_kernel_and_agree_with_the_oracle():
    """RGBA rows whose width is not a multiple of 4 (neither input rows nor plane rows are 16-byte
    aligned): planar output goes through k1_window (variant 5); same bits as the oracle, as one pixel per
    thread and as the scalar kernel, for flips, dropped alpha, every output type and images whose plane
    size leaves every alignment phase (H*W mod 4 = 1, 2, 3) and whose windows straddle rows."""
    from blendtorch.btt import native as N
    rng = np.random.default_rng(11)
    for shape in ((3, 37, 101, 4), (2, 5, 3, 4), (4, 9, 6, 4), (5, 1, 7, 4), (2, 3, 1, 4), (1, 1, 1, 4), (3, 13, 1022, 4)):
        frames = rng.integers(0, 256, shape, dtype=np.uint8)
        for dtype, flip, oc in itertools.product((torch.float32, torch.bfloat16, torch.uint8), (False, True), (None, 3)):
            kw = dict(flip=flip, gamma=2.2, dtype=dtype, out_channels=oc)
            if dtype != torch.uint8:
                kw.update(mean=[100.0, 110.0, 120.0, 130.0], std=[50.0, 60.0, 70.0, 80.0])
            got = _run(frames, **kw)
            assert N.lib.bt_preprocess_last_variant() == 5
            _assert_matches(got, _oracle(frames, flip, 2.2, kw.get("mean"), kw.get("std"), "NCHW", dtype, oc), dtype)
            px = _run(frames, variant="pixel", **kw)
            assert N.lib.bt_preprocess_last_variant() == 4
            assert torch.equal(got.view(torch.uint8), px.view(torch.uint8))
            assert torch.equal(got.view(torch.uint8), _run(frames, variant="generic", **kw).view(torch.uint8))


def test_stats_match_oracle_and_are_deterministic():
    rng = np.random.default_rng(9)
    frames = rng.integers(0, 256, (6, 40, 64, 4), dtype=np.uint8)
    for kw in (dict(gamma=2.2, mean=127.5, std=127.5), dict(dtype=torch.uint8, layout="NHWC"),
               dict(mean=[1., 2., 3., 4.], std=[5., 6., 7., 8.], dtype=torch.bfloat16, out_channels=3)):
        out, st = _run(frames, stats=True, **kw)
        s, q = opre.batch_stats(out.float().cpu().numpy() if kw.get("layout") != "NHWC"
                                else out.float().permute(0, 3, 1, 2).cpu().numpy())
        got = st.cpu().numpy()
        np.testing.assert_allclose(got[:, 0], s, rtol=1e-6, atol=1e-3)   # fp32 partial sums of <=64 values
        np.testing.assert_allclose(got[:, 1], q, rtol=1e-6, atol=1e-3)
        out2, st2 = _run(frames, stats=True, **kw)
        assert torch.equal(st, st2)                                        # fixed reduction order
    # generic variant too (every value goes straight into a float64 sum)
    out, st = _run(frames[:, :, :5], stats=True, mean=127.5, std=127.5, variant="generic")
    s, q = opre.batch_stats(out.cpu().numpy())
    np.testing.assert_allclose(st.cpu().numpy()[:, 0], s, rtol=1e-9, atol=1e-6)
    # odd widths: the aligned-group kernel (fp32 partial sums of 4 values) and one pixel per thread
    for variant in ("auto", "pixel"):
        out, st = _run(frames[:, :, :5], stats=True, mean=127.5, std=127.5, variant=variant)
        s, q = opre.batch_stats(out.cpu().numpy())
        np.testing.assert_allclose(st.cpu().numpy()[:, 0], s, rtol=1e-6, atol=1e-3)
        np.testing.assert_allclose(st.cpu().numpy()[:, 1], q, rtol=1e-6, atol=1e-3)


def test_empty_and_strided_batches():
    out = _run(np.zeros((0, 4, 8, 4), np.uint8), mean=127.5, std=127.5)
    assert out.shape == (0, 4, 4, 8)
    # frames that sit inside a wider staging buffer (image stride > dense size)
    rng = np.random.default_rng(1)
    big = torch.from_numpy(rng.integers(0, 256, (3, 4 * 8 * 4 + 256), dtype=np.uint8)).to(DEV)
    view = big[:, :4 * 8 * 4].view(3, 4, 8, 4)
    out = btt.preprocess_images(view, btt.Preprocess(mean=127.5, std=127.5))
    exp = opre.preprocess_batch(list(view.cpu().numpy()), mean=127.5, std=127.5)
    np.testing.assert_array_equal(out.cpu().numpy(), exp)


def _lut_expected(src, lut_np, flip, chw):
    """Full-size check: gather through the oracle's 256-entry table with torch ops on the GPU."""
    B, H, W, C = src.shape
    lut = torch.from_numpy(lut_np).to(src.device)          # (C, 256)
    idx = src.long()
    if flip:
        idx = idx.flip(1)
    out = torch.stack([lut[c][idx[..., c]] for c in range(C)], dim=1 if chw else -1)
    return out


@pytest.mark.parametrize("name,shape", [("C2", (8, 480, 640, 4)), ("C3", (32, 1080, 1920, 4)),
                                        ("C4", (64, 1024, 1024, 4)), ("C5", (1024, 128, 128, 4))])
def test_full_size_properties(name, shape):
    """BASELINE.json sizes: (a) exactness against the oracle's per-byte table gathered by torch,
    (b) flip is an involution, (c) the fused statistics equal torch's own reduction."""
    g = torch.Generator(device=DEV)
    g.manual_seed(1234)
    src = torch.randint(0, 256, shape, dtype=torch.uint8, device=DEV, generator=g)
    ramp = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(4, axis=2)
    lut = opre.preprocess_frame(ramp, gamma=2.2, mean=127.5, std=127.5)[:, 0, :]   # oracle on all 256 bytes
    spec = btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True)
    out, st = btt.preprocess_images(src, spec)
    exp = _lut_expected(src, lut, True, True)
    assert torch.equal(out, exp)
    # statistics accumulate fp32 partial sums of <= 16 values into float64: error <= 2^-22 of sum|x|
    tot = out.double().sum(dim=(0, 2, 3))
    scale = out.double().abs().sum(dim=(0, 2, 3))
    assert ((st[:, 0] - tot).abs() <= 2.0 ** -22 * scale).all()
    sq = (out.double() ** 2).sum(dim=(0, 2, 3))
    assert ((st[:, 1] - sq).abs() <= 2.0 ** -22 * sq).all()
    del exp, out
    u8 = btt.Preprocess(flip=True, dtype=torch.uint8, layout="NHWC")
    twice = btt.preprocess_images(btt.preprocess_images(src, u8), u8)
    assert torch.equal(twice, src)
    bf = btt.preprocess_images(src, btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5, dtype=torch.bfloat16))
    exp = _lut_expected(src, lut, True, True).to(torch.bfloat16)
    assert torch.equal(bf, exp)


def test_bulk_async_variant_matches_the_register_variant():
    """variant='bulk' (cp.async.bulk + mbarrier ring staging the input in shared memory) is the measured
    A/B partner of the register-vectorised kernel: identical bits, including ragged last segments
    (W*4 not a multiple of the 4 KiB segment), flips, alpha dropped, every output type."""
    from blendtorch.btt import native as N
    rng = np.random.default_rng(21)
    for shape in ((3, 9, 1028, 4), (2, 5, 640, 4), (1, 3, 4, 4), (2, 17, 2052, 4), (40, 33, 64, 4)):
        frames = rng.integers(0, 256, shape, dtype=np.uint8)
        for dtype in (torch.float32, torch.bfloat16, torch.uint8):
            for flip in (False, True):
                kw = dict(flip=flip, gamma=2.2, dtype=dtype)
                if dtype != torch.uint8:
                    kw.update(mean=127.5, std=127.5)
                a = _run(frames, variant="vector", **kw)
                b = _run(frames, variant="bulk", **kw)
                assert N.lib.bt_preprocess_last_variant() == 3
                assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
    out = _run(frames, variant="bulk", gamma=2.2, mean=127.5, std=127.5, out_channels=3)
    _assert_matches(out, _oracle(frames, False, 2.2, 127.5, 127.5, "NCHW", torch.float32, 3), torch.float32)
    with pytest.raises(N.NativeError):
        _run(rng.integers(0, 256, (1, 4, 8, 3), dtype=np.uint8), variant="bulk")
    with pytest.raises(N.NativeError):
        _run(frames, variant="bulk", layout="NHWC")


def test_gather_index_matches_oracle_on_every_variant():
    """`index=` draws the batch from frames resident on the device: output item i must be the oracle's
    transform of frame index[i] (repeats, reversed order, a subset), bit for bit, on every kernel."""
    from blendtorch.btt import native as N
    rng = np.random.default_rng(33)
    cases = [((9, 12, 16, 4), "auto", 2), ((9, 12, 16, 4), "generic", 1), ((6, 8, 1028, 4), "bulk", 3),
             ((7, 7, 5, 4), "auto", None), ((7, 7, 5, 4), "pixel", 4), ((5, 9, 16, 3), "auto", None), ((4, 5, 16, 1), "auto", None)]
    for shape, variant, ran in cases:
        frames = rng.integers(0, 256, shape, dtype=np.uint8)
        src = torch.from_numpy(frames).to(DEV)
        idx = np.concatenate([rng.permutation(shape[0]), [0, 0, shape[0] - 1]])
        index = torch.from_numpy(idx).to(DEV)
        for dtype, layout, flip in ((torch.float32, "NCHW", True), (torch.bfloat16, "NCHW", False),
                                    (torch.float32, "NHWC", True), (torch.uint8, "NHWC", False)):
            if variant == "bulk" and layout == "NHWC":
                continue
            kw = dict(flip=flip, gamma=2.2, layout=layout, dtype=dtype, variant=variant)
            if dtype != torch.uint8:
                kw.update(mean=127.5, std=127.5)
            out = btt.preprocess_images(src, btt.Preprocess(**kw), index=index)
            torch.cuda.synchronize()
            if ran is not None:
                assert N.lib.bt_preprocess_last_variant() == ran, (shape, variant, layout)
            elif shape == (7, 7, 5, 4):     # odd width: planar output by aligned groups, interleaved by pixels
                assert N.lib.bt_preprocess_last_variant() == (5 if layout == "NCHW" else 4)
            assert out.shape[0] == len(idx)
            exp = _oracle(frames[idx], flip, 2.2, kw.get("mean"), kw.get("std"), layout, dtype)
            _assert_matches(out, exp, dtype)
    # statistics follow the gathered items, and an empty index gives an empty batch
    out, st = btt.preprocess_images(src, btt.Preprocess(mean=127.5, std=127.5, stats=True), index=index)
    dense, st2 = btt.preprocess_images(src[index], btt.Preprocess(mean=127.5, std=127.5, stats=True))
    assert torch.equal(out, dense) and torch.equal(st, st2)
    none = btt.preprocess_images(src, btt.Preprocess(), index=index[:0])
    assert none.shape[0] == 0
    with pytest.raises(ValueError):
        btt.preprocess_images(src, btt.Preprocess(), index=index.to(torch.int32))


def test_c_program_drives_the_abi_without_python(tmp_path):
    """examples/c_abi_demo.c: cudaMalloc + bt_preprocess_launch from plain C, checked in C against the
    reference's expression (exit status 0 = every value bit-identical)."""
    import subprocess
    from test_abi_and_rules import _build_c_demo
    exe = _build_c_demo(tmp_path / "c_abi_demo")
    res = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "0 of" in res.stdout


def test_flip_rows_any_dtype_and_gather():
    """bt_flip_rows_launch == torch.flip(dims=[1]) for float32 depth maps, odd row sizes (4- and 1-byte
    units), uint16 masks, and with a gather index."""
    g = torch.Generator(device=DEV)
    g.manual_seed(3)
    cases = [((5, 9, 16), torch.float32), ((3, 7, 5), torch.float32), ((2, 6, 7), torch.uint8), ((4, 5, 3, 2), torch.int16),
             ((2, 1080, 1920), torch.float32), ((3, 1, 8), torch.float64), ((0, 4, 4), torch.float32)]
    for shape, dt in cases:
        src = (torch.rand(shape, device=DEV, generator=g) * 1000).to(dt)
        out = btt.flip_rows(src)
        assert torch.equal(out, torch.flip(src, dims=[1])), (shape, dt)
        if shape[0] > 1:
            idx = torch.tensor([shape[0] - 1, 0, 0], device=DEV)
            assert torch.equal(btt.flip_rows(src, index=idx), torch.flip(src[idx], dims=[1]))
    with pytest.raises(ValueError):
        btt.flip_rows(torch.zeros(4, device=DEV))
