#!/usr/bin/env python3
"""Generate the golden fixtures in this directory by running the UNMODIFIED
reference (/root/reference) in the build container.

Run once, by hand, where /root/reference exists:

    python tests/golden/make_golden.py

Nothing in the test-suite, smoke() or bench.py executes this file or reads
/root/reference; they only read the small files it leaves behind.  The
reference's btb modules `import bpy` (Blender) and the GL bindings at module
top; those imports are satisfied with empty stand-in modules so that the
*numpy* lines of the reference (flip, gamma, projection) can be executed as
they are.  No graphics API is touched.

What is produced (all deterministic):
  rec_small_00.btr, rec_small_01.btr   recordings written by the reference's own
        FileRecorder.save(pickle.dumps(msg, 3), is_pickled=True) -- the path
        btt/dataset.py:100-103 takes
  rec_small_expected.npz    the same recordings replayed through the reference's
        FileDataset + DataLoader(default_collate): raw, with generate.py's
        gamma_correct, with densityopt.py's normalise+CHW, and with
        offscreen.py's flip+_color_correct
  file_roundtrip.btr        the exact protocol of tests/test_file.py:5-13
  gamma_tables.npz          offscreen._color_correct / generate.gamma_correct over
        all 256 byte values for several coefficients
  camera_golden.npz         Camera.world_to_ndc / ndc_to_pixel (reference code)
        on the default cube and on random points/matrices, plus the expected
        arrays of tests/test_camera.py:19-40
"""

import ast
import pickle
import sys
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
HERE = Path(__file__).resolve().parent


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def import_reference():
    class _Vector(list):
        pass

    _stub("bpy")
    _stub("bpy_extras")
    _stub("bmesh")
    _stub("mathutils", Vector=_Vector)
    _stub("gpu")
    _stub("bgl")
    _stub("OpenGL")
    _stub("OpenGL.GL", glGetTexImage=None)
    sys.path.insert(0, str(REF / "pkg_pytorch"))
    sys.path.insert(0, str(REF / "pkg_blender"))
    from blendtorch import btt  # noqa
    import importlib

    cam = importlib.import_module("blendtorch.btb.camera")
    off = importlib.import_module("blendtorch.btb.offscreen")
    return btt, cam, off


def extract_function(pyfile, name):
    """Compile one top-level function of a reference example without running
    the module (the examples import matplotlib / launch Blender at top level)."""
    src = Path(pyfile).read_text()
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            mod = ast.Module(body=[node], type_ignores=[])
            ns = {"np": np}
            exec(compile(mod, str(pyfile), "exec"), ns)
            return ns[name]
    raise KeyError(name)


def main():
    btt, refcam, refoff = import_reference()
    from torch.utils.data import DataLoader

    sys.path.insert(0, str(HERE.parent.parent))
    from oracle import synth  # inputs only

    gamma_correct = extract_function(REF / "examples/datagen/generate.py", "gamma_correct")
    densityopt_transform = extract_function(REF / "examples/densityopt/densityopt.py", "item_transform")

    # ---- recordings through the reference recorder -------------------------------
    H, W = 12, 16
    prefix = HERE / "rec_small"
    counts = (6, 5)
    sent = []
    for pub, n in enumerate(counts):
        rng = synth.rng_for(1, pub)
        with btt.FileRecorder(btt.FileRecorder.filename(prefix, pub), max_messages=8) as rec:
            for i in range(n):
                msg = synth.message(rng, 1, pub, i, H, W)
                if i == 2:
                    msg["image"] = synth.gradient_frame(H, W)
                if i == 3:
                    msg["image"] = np.full((H, W, 4), 255, np.uint8)
                if i == 4:
                    msg["image"] = np.zeros((H, W, 4), np.uint8)
                sent.append(msg)
                rec.save(pickle.dumps(msg, protocol=3), is_pickled=True)

    def replay(transform, batch):
        ds = btt.FileDataset(str(prefix), item_transform=transform)  # fresh per loader (defect 4)
        return list(DataLoader(ds, batch_size=batch, shuffle=False, num_workers=0))

    out = {}
    raw = replay(None, 4)
    for k, b in enumerate(raw):
        for key in ("image", "xy", "frameid", "btid"):
            out["raw_b%d_%s" % (k, key)] = b[key].numpy()

    def t_gamma(item):
        item["image"] = gamma_correct(item["image"])
        return item

    for k, b in enumerate(replay(t_gamma, 4)):
        out["gamma_b%d_image" % k] = b["image"].numpy()

    def t_norm(item):
        x, sid = densityopt_transform({"image": item["image"], "shape_id": item["frameid"]})
        return {"image": x, "shape_id": sid}

    for k, b in enumerate(replay(t_norm, 4)):
        out["norm_b%d_image" % k] = b["image"].numpy()
        out["norm_b%d_shape_id" % k] = b["shape_id"].numpy()

    def t_gamma_norm(item):
        item["image"] = gamma_correct(item["image"])
        return t_norm(item)

    for k, b in enumerate(replay(t_gamma_norm, 4)):
        out["gamma_norm_b%d_image" % k] = b["image"].numpy()

    # producer tail of offscreen.py:94-99 applied to a bottom-up frame: flip then gamma
    def t_offscreen(item):
        buf = np.flipud(item["image"])                       # offscreen.py:95-96
        buf = refoff.OffScreenRenderer._color_correct(None, buf, 2.2)  # :97-98, :105-112
        item["image"] = np.ascontiguousarray(buf)
        return item

    for k, b in enumerate(replay(t_offscreen, 4)):
        out["offscreen_b%d_image" % k] = b["image"].numpy()

    def t_offscreen_rgb(item):
        buf = np.flipud(item["image"][..., :3])
        item["image"] = np.ascontiguousarray(refoff.OffScreenRenderer._color_correct(None, buf, 2.2))
        return item

    for k, b in enumerate(replay(t_offscreen_rgb, 4)):
        out["offscreen_rgb_b%d_image" % k] = b["image"].numpy()

    np.savez_compressed(HERE / "rec_small_expected.npz", **out)

    # ---- tests/test_file.py protocol ---------------------------------------------
    with btt.FileRecorder(outpath=HERE / "file_roundtrip.btr", max_messages=10) as rec:
        for i in range(7):
            rec.save({"value": i}, is_pickled=False)
    r = btt.FileReader(HERE / "file_roundtrip.btr")
    assert len(r) == 7 and [r[i]["value"] for i in range(7)] == list(range(7))

    # ---- gamma tables ------------------------------------------------------------
    ramp = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(4, axis=2)
    tabs = {}
    for g in (1.0, 1.8, 2.0, 2.2, 2.4, 0.5):
        a = refoff.OffScreenRenderer._color_correct(None, ramp, g)
        tabs["g%s" % str(g).replace(".", "_")] = a[0, :, 0].copy()
        assert (a[..., 3] == ramp[..., 3]).all()
    b = gamma_correct(ramp)
    assert (b[0, :, 0] == tabs["g2_2"]).all()
    np.savez_compressed(HERE / "gamma_tables.npz", **tabs)

    # ---- camera -------------------------------------------------------------------
    from oracle import camera as ocam

    def ref_camera(V, P, shape):
        c = object.__new__(refcam.Camera)
        c.view_matrix, c.proj_matrix, c.shape = V, P, shape
        return c

    shape = (480, 640)
    cube = ocam.default_cube_vertices()
    V = ocam.view_looking_down_z(7.0)
    cg = {
        "cube": cube, "V": V,
        "P_persp": ocam.blender_perspective(shape),
        "P_ortho": ocam.blender_orthographic(shape),
        # tests/test_camera.py:19-40 (golden numbers of the reference's own test)
        "expected_ortho_xy": np.array([[480., 80], [480, 80], [480, 400], [480, 400],
                                       [160, 80], [160, 80], [160, 400], [160, 400]]),
        "expected_proj_xy": np.array([[468.148, 91.851], [431.111, 128.888], [468.148, 388.148],
                                      [431.111, 351.111], [171.851, 91.851], [208.888, 128.888],
                                      [171.851, 388.148], [208.888, 351.111]]),
        "expected_z": np.array([6., 8, 6, 8, 6, 8, 6, 8]),
    }
    for tag in ("persp", "ortho"):
        c = ref_camera(np.asarray(V), np.asarray(cg["P_" + tag]), shape)
        ndc, z = c.world_to_ndc(cube, return_depth=True)
        cg["ref_%s_ndc" % tag] = ndc
        cg["ref_%s_z" % tag] = z
        cg["ref_%s_px_ul" % tag] = c.ndc_to_pixel(ndc.copy(), origin="upper-left")
        cg["ref_%s_px_ll" % tag] = c.ndc_to_pixel(ndc.copy(), origin="lower-left")
        cg["ref_%s_ndc_nodepth" % tag] = c.world_to_ndc(cube)
    rng = np.random.default_rng(3000)
    pts = synth.cube_corners(rng, 8)
    ang = 0.3
    Vr = np.array([[np.cos(ang), 0, np.sin(ang), 0.5], [0, 1, 0, -0.25],
                   [-np.sin(ang), 0, np.cos(ang), -2.0], [0, 0, 0, 1]])
    shape2 = (1080, 1920)
    Pr = ocam.blender_perspective(shape2, lens=35.0)
    c = ref_camera(Vr, Pr, shape2)
    ndc, z = c.world_to_ndc(pts, return_depth=True)
    cg.update(rand_pts=pts, rand_V=Vr, rand_P=Pr, rand_ndc=ndc, rand_z=z,
              rand_px_ul=c.ndc_to_pixel(ndc.copy(), origin="upper-left"),
              rand_px_ll=c.ndc_to_pixel(ndc.copy(), origin="lower-left"),
              rand_ndc_nodepth=c.world_to_ndc(pts))
    np.savez_compressed(HERE / "camera_golden.npz", **cg)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
