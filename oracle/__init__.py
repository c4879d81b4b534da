"""CPU oracle for the blendtorch btt-side ingest path.  TEST INFRASTRUCTURE ONLY.

This package restates, in plain numpy / stdlib pickle, what the reference
(v-wewei/pytorch-blender, "blendtorch" 0.2.0) computes on the path

    pickled ZMQ / .btr message -> dict of ndarrays -> item_transform
        -> default_collate -> batch of tensors

so that the CUDA path can be compared against it byte for byte.  Each function
cites the reference file:line it follows (paths relative to the reference root).

Rules (enforced by tests/test_abi_and_rules.py):
  * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
    --impl reference legs may import anything from here;
  * the product package (pytorch-blender_b200/) never imports it and has no CPU
    fallback for its kernels.

Pinning: the restatement is pinned against (a) the reference's own golden
vectors (tests/test_camera.py:19-40, tests/test_file.py:5-13) and (b) outputs of
the unmodified reference executed in the build container, committed as fixtures
under tests/golden/ together with the script that made them
(tests/golden/make_golden.py).  The reference is pure Python, so there is no
oracle/_ref binary; `oracle/_ref/` stays empty by design.

Third-party arithmetic the reference delegates to and that is therefore part of
the oracle by call, not by restatement: CPython `pickle` (protocol 3 on the
Blender side, btt/file.py:63,72), numpy ndarray pickling, and
`torch.utils.data.default_collate` (installed torch 2.11, collate.py:118-313).
"""

from . import btr, camera, preprocess, raster, stream, synth  # noqa: F401
