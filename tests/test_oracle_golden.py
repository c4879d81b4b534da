"""The oracle restatement pinned against the reference: its own golden vectors
(tests/test_camera.py:19-40, tests/test_file.py:5-13) and the outputs of the
unmodified reference recorded by tests/golden/make_golden.py."""
import pickle

import numpy as np
import pytest
from numpy.testing import assert_allclose, assert_array_equal

from oracle import btr, camera, preprocess, synth


@pytest.fixture(scope="module")
def expected(golden_dir):
    return dict(np.load(golden_dir / "rec_small_expected.npz"))


def _batches(golden_dir, transform, batch=4):
    ds = btr.ReplayDataset(str(golden_dir / "rec_small"), item_transform=transform)
    items = [ds[i] for i in range(len(ds))]
    return [preprocess.collate(items[i:i + batch]) for i in range(0, len(items), batch)]


def test_reader_matches_reference_replay(golden_dir, expected):
    bs = _batches(golden_dir, None)
    assert len(bs) == 3
    for k, b in enumerate(bs):
        for key in ("image", "xy", "frameid", "btid"):
            got = b[key].numpy()
            exp = expected["raw_b%d_%s" % (k, key)]
            assert got.dtype == exp.dtype and got.shape == exp.shape
            assert_array_equal(got, exp)
    assert bs[0]["image"].numpy().dtype == np.uint8
    assert bs[0]["xy"].numpy().dtype == np.float64
    assert bs[0]["frameid"].numpy().dtype == np.int64


def test_recorder_bytes_identical_to_reference(golden_dir, tmp_path):
    """Re-record the same messages with the oracle recorder: files must be
    byte-identical to the ones the reference's FileRecorder wrote."""
    for pub, n in enumerate((6, 5)):
        src = btr.Reader(golden_dir / ("rec_small_%02d.btr" % pub))
        assert len(src) == n
        with btr.Recorder(tmp_path / ("x_%02d.btr" % pub), max_messages=8) as rec:
            for i in range(n):
                rec.save(pickle.dumps(src[i], protocol=3), is_pickled=True)
        a = (golden_dir / ("rec_small_%02d.btr" % pub)).read_bytes()
        b = (tmp_path / ("x_%02d.btr" % pub)).read_bytes()
        assert a == b


def test_file_roundtrip_fixture(golden_dir):
    r = btr.Reader(golden_dir / "file_roundtrip.btr")  # tests/test_file.py:5-13
    assert len(r) == 7
    for i in range(7):
        assert r[i]["value"] == i


def test_gamma_matches_reference(golden_dir, expected):
    def t(item):
        item["image"] = preprocess.gamma_correct(item["image"])
        return item
    for k, b in enumerate(_batches(golden_dir, t)):
        assert_array_equal(b["image"].numpy(), expected["gamma_b%d_image" % k])


def test_gamma_tables(golden_dir):
    tabs = np.load(golden_dir / "gamma_tables.npz")
    for name in tabs.files:
        g = float(name[1:].replace("_", "."))
        assert_array_equal(preprocess.gamma_table(g), tabs[name])
    # alpha is passed through, colour goes through the table
    rng = np.random.default_rng(0)
    f = rng.integers(0, 256, (5, 7, 4), dtype=np.uint8)
    out = preprocess.gamma_correct(f, 2.2)
    assert_array_equal(out[..., 3], f[..., 3])
    assert_array_equal(out[..., :3], tabs["g2_2"][f[..., :3]])


def test_normalise_matches_reference(golden_dir, expected):
    def t(item):
        return {"image": preprocess.normalise_chw(item["image"]), "shape_id": item["frameid"]}
    for k, b in enumerate(_batches(golden_dir, t)):
        got = b["image"].numpy()
        assert got.dtype == np.float32
        assert_array_equal(got, expected["norm_b%d_image" % k])
        assert_array_equal(b["shape_id"].numpy(), expected["norm_b%d_shape_id" % k])


def test_composed_preprocess_matches_reference(golden_dir, expected):
    ds = btr.ReplayDataset(str(golden_dir / "rec_small"))
    frames = [ds[i]["image"] for i in range(len(ds))]
    got = preprocess.preprocess_batch(frames, gamma=2.2, mean=127.5, std=127.5)
    exp = np.concatenate([expected["gamma_norm_b%d_image" % k] for k in range(3)])
    assert_array_equal(got, exp)
    got = preprocess.preprocess_batch(frames, flip=True, gamma=2.2, chw=False, out_dtype="uint8")
    exp = np.concatenate([expected["offscreen_b%d_image" % k] for k in range(3)])
    assert_array_equal(got, exp)
    got = preprocess.preprocess_batch([f[..., :3] for f in frames], flip=True, gamma=2.2,
                                      chw=False, out_dtype="uint8")
    exp = np.concatenate([expected["offscreen_rgb_b%d_image" % k] for k in range(3)])
    assert_array_equal(got, exp)


def test_camera_reference_goldens(golden_dir):
    g = np.load(golden_dir / "camera_golden.npz")
    shape = (480, 640)
    for tag, exp_xy in (("ortho", g["expected_ortho_xy"]), ("persp", g["expected_proj_xy"])):
        px, z = camera.world_to_pixel(g["cube"], g["V"], g["P_" + tag], shape, return_depth=True)
        assert_allclose(px, exp_xy, atol=1e-2)          # tests/test_camera.py:46-49
        assert_allclose(z, g["expected_z"], atol=1e-2)
        ndc, z2 = camera.world_to_ndc(g["cube"], g["V"], g["P_" + tag], True)
        assert_array_equal(ndc, g["ref_%s_ndc" % tag])   # bit-exact vs reference code
        assert_array_equal(z2, g["ref_%s_z" % tag])
        assert_array_equal(camera.ndc_to_pixel(ndc.copy(), shape, "upper-left"), g["ref_%s_px_ul" % tag])
        assert_array_equal(camera.ndc_to_pixel(ndc.copy(), shape, "lower-left"), g["ref_%s_px_ll" % tag])
        assert_array_equal(camera.world_to_ndc(g["cube"], g["V"], g["P_" + tag]), g["ref_%s_ndc_nodepth" % tag])
    ndc, z = camera.world_to_ndc(g["rand_pts"], g["rand_V"], g["rand_P"], True)
    assert_array_equal(ndc, g["rand_ndc"])
    assert_array_equal(z, g["rand_z"])
    assert_array_equal(camera.ndc_to_pixel(ndc.copy(), (1080, 1920)), g["rand_px_ul"])
    assert_array_equal(camera.world_to_ndc(g["rand_pts"], g["rand_V"], g["rand_P"]), g["rand_ndc_nodepth"])


def test_pixel_bboxes():
    rng = np.random.default_rng(1)
    xy = rng.random((24, 2)) * 100
    bb = camera.pixel_bboxes(xy)
    assert bb.shape == (3, 4)
    for m in range(3):
        c = xy[m * 8:(m + 1) * 8]
        assert_array_equal(bb[m], [c[:, 0].min(), c[:, 1].min(), c[:, 0].max(), c[:, 1].max()])


def test_bf16_bits_match_torch():
    import torch
    x = np.random.default_rng(2).standard_normal(4096).astype(np.float32)
    exp = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert_array_equal(preprocess.to_bf16_bits(x), exp)
