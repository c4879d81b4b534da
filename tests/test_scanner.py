"""Message framing (bt_scan_message / bt_schema_*) against stdlib pickle on the host.
Replaces pickle.loads at btt/dataset.py:103,105 and btt/file.py:108."""
import pickle

import numpy as np
import pytest

from blendtorch.btt import native as N
from blendtorch.btt.staging import Ring, Schema


def _msg(**extra):
    rng = np.random.default_rng(7)
    m = {"btid": 3, "image": rng.integers(0, 256, (6, 10, 4), dtype=np.uint8),
         "xy": rng.random((8, 2)), "frameid": 41}
    m.update(extra)
    return m


def _check_fields(blob, obj):
    fields = {f.key.decode(): f for f in N.scan_message(blob)}
    assert list(fields) == list(obj)
    for k, v in obj.items():
        f = fields[k]
        if isinstance(v, (np.ndarray, np.generic)):
            a = np.asarray(v)
            assert f.kind == N.BT_KIND_NDARRAY
            assert N.NUMPY_DTYPES[f.dtype] == a.dtype.name
            assert [f.shape[d] for d in range(f.ndim)] == list(a.shape)
            order = "F" if f.fortran else "C"
            assert bytes(blob[f.payload_off:f.payload_off + f.payload_len]) == a.tobytes(order=order)
        elif isinstance(v, bool):
            assert f.kind == N.BT_KIND_BOOL and bool(f.ival) == v
        elif isinstance(v, int):
            assert f.kind == N.BT_KIND_INT and f.ival == v
        elif isinstance(v, float):
            assert f.kind == N.BT_KIND_FLOAT and f.fval == v
        elif isinstance(v, str):
            assert f.kind == N.BT_KIND_STR
            assert bytes(blob[f.payload_off:f.payload_off + f.payload_len]).decode() == v
        elif v is None:
            assert f.kind == N.BT_KIND_NONE
        else:
            assert f.kind == N.BT_KIND_OTHER


@pytest.mark.parametrize("protocol", [2, 3, 4, 5])
def test_scan_matches_pickle(protocol):
    obj = _msg(f=2.5, flag=True, neg=-5, big=2 ** 40, name="cube", nothing=None,
               depth=np.linspace(0, 1, 12, dtype=np.float32).reshape(3, 4),
               idx=np.arange(5, dtype=np.int64), scalar=np.float64(1.25), nested=[1, 2, (3, 4)])
    if protocol == 2:
        # protocol 2 has no bytes opcode: numpy sends latin-1 text through _codecs.encode, which
        # the scanner records but never calls -> arrays are reported as OTHER and the schema
        # refuses them, so the caller falls back to pickle.loads
        blob = pickle.dumps(obj, protocol=2)
        kinds = {f.key.decode(): f.kind for f in N.scan_message(blob)}
        assert kinds["image"] == N.BT_KIND_OTHER and kinds["frameid"] == N.BT_KIND_INT
        sch = Schema.from_message(blob, 2)
        assert not sch.native and sch.column("image").shape == (6, 10, 4)
        return
    blob = pickle.dumps(obj, protocol=protocol)
    _check_fields(blob, obj)


@pytest.mark.parametrize("protocol", [3, 4])
def test_numpy1_module_path(protocol):
    """Blender-era numpy 1.x pickles name `numpy.core.multiarray`; numpy 2 writes `numpy._core`."""
    obj = _msg()
    blob = pickle.dumps(obj, protocol=protocol)
    if protocol == 3:
        old = blob.replace(b"numpy._core.multiarray\n", b"numpy.core.multiarray\n")
    else:
        name = b"numpy._core.multiarray"
        i = blob.index(name)
        assert blob[i - 1] == len(name)
        old = blob[:i - 1] + bytes([len(name) - 1]) + b"numpy.core.multiarray" + blob[i + len(name):]
        # the enclosing FRAME length is not checked by the scanner (nor needed by it)
    assert old != blob
    _check_fields(old, obj)


def test_shapes_dtypes_and_edge_cases():
    for dt in ("uint8", "int8", "uint16", "int16", "uint32", "int32", "uint64", "int64",
               "float16", "float32", "float64", "bool"):
        for shape in ((), (0,), (1,), (3, 0, 2), (2, 3, 4, 5, 2)):
            a = np.zeros(shape, dtype=dt)
            if a.size:
                a.reshape(-1)[:] = np.arange(a.size) % 2
            obj = {"a": a}
            for proto in (3, 4, 5):
                _check_fields(pickle.dumps(obj, protocol=proto), obj)
    # non-contiguous and flipped views pickle as their logical C-order bytes
    base = np.arange(4 * 5 * 3, dtype=np.uint8).reshape(4, 5, 3)
    for view in (base[::-1], base[:, ::2], np.flipud(base)):
        obj = {"image": view}
        _check_fields(pickle.dumps(obj, protocol=3), obj)
    # Fortran order is reported, not silently misread
    f = np.asfortranarray(np.arange(12, dtype=np.float32).reshape(3, 4))
    fld = N.scan_message(pickle.dumps({"f": f}, protocol=3))[0]
    assert fld.fortran == 1
    # same array under two keys (memo BINGET)
    a = np.arange(6, dtype=np.int32)
    obj = {"a": a, "b": a}
    _check_fields(pickle.dumps(obj, protocol=3), obj)
    # big-endian payload -> not a column
    be = np.arange(4, dtype=">i4")
    assert N.scan_message(pickle.dumps({"x": be}, protocol=3))[0].kind == N.BT_KIND_OTHER
    # huge python int
    assert N.scan_message(pickle.dumps({"x": 2 ** 70}, protocol=3))[0].kind == N.BT_KIND_OTHER
    assert N.scan_message(pickle.dumps({"x": -2 ** 63}, protocol=3))[0].ival == -2 ** 63


def test_truncated_and_hostile_input():
    blob = pickle.dumps(_msg(), protocol=3)
    for cut in (0, 1, 5, 50, 200, len(blob) - 1):
        with pytest.raises(N.NativeError) as e:
            N.scan_message(blob[:cut])
        assert e.value.code in (N.BT_E_TRUNCATED, N.BT_E_UNSUPPORTED)
    with pytest.raises(N.NativeError):
        N.scan_message(pickle.dumps([1, 2, 3], protocol=3))      # not a dict
    with pytest.raises(N.NativeError):
        N.scan_message(b"\x80\x03cos\nsystem\nX\x02\x00\x00\x00idq\x00\x85R.")  # REDUCE is recorded, never called
    rng = np.random.default_rng(0)
    for _ in range(300):  # fuzz: must never crash, only report
        junk = rng.integers(0, 256, rng.integers(1, 200), dtype=np.uint8).tobytes()
        try:
            N.scan_message(junk)
        except N.NativeError:
            pass
    for _ in range(300):  # mutated valid messages
        b = bytearray(blob[:400])
        for _ in range(3):
            b[rng.integers(0, len(b))] = rng.integers(0, 256)
        try:
            N.scan_message(bytes(b) + blob[400:])
        except N.NativeError:
            pass


def test_schema_layout_and_ingest_into_unpinned_ring():
    B = 5
    msgs = []
    rng = np.random.default_rng(11)
    for i in range(B):
        msgs.append({"btid": i % 2, "image": rng.integers(0, 256, (6, 10, 4), dtype=np.uint8),
                     "xy": rng.random((8, 2)), "frameid": 100 + i, "score": 0.5 * i, "ok": bool(i & 1),
                     "depth": rng.random((6, 10), dtype=np.float32)})
    blobs = [pickle.dumps(m, protocol=3 + (i % 2)) for i, m in enumerate(msgs)]
    sch = Schema.from_message(blobs[0], B)
    assert sch.native
    assert [c.key for c in sch.columns] == list(msgs[0])
    for c in sch.columns:
        assert c.slot_off % 256 == 0
    ring = Ring(sch.slot_bytes, depth=2, device=None)
    slot = ring.acquire()
    for i, b in enumerate(blobs):
        sch.ingest(ring.ptr(slot), ring.view(slot), i, b)
    view = ring.view(slot)
    from torch.utils.data import default_collate
    exp = default_collate(msgs)
    for c in sch.columns:
        got = view[c.slot_off:c.slot_off + B * c.item_bytes].view(c.dtype).reshape((B,) + c.shape)
        e = exp[c.key].numpy()
        assert got.dtype == e.dtype, c.key      # collated dtypes are part of the contract
        np.testing.assert_array_equal(got, e)
    # bulk, multi-threaded variant gives the same bytes
    slot2 = ring.acquire()
    addrs = np.array([N.buffer_address(b)[0] for b in blobs], dtype=np.uint64)
    lens = np.array([len(b) for b in blobs], dtype=np.uint64)
    assert sch.ingest_many(ring.ptr(slot2), 0, addrs, lens, nthreads=3) == -1
    np.testing.assert_array_equal(ring.view(slot2)[:sch.slot_bytes], view[:sch.slot_bytes])
    # a message whose image changed shape is reported, not copied
    bad = pickle.dumps({**msgs[0], "image": np.zeros((6, 11, 4), np.uint8)}, protocol=3)
    rc = N.lib.bt_schema_ingest(sch.handle, ring.ptr(slot), 0, *N.buffer_address(bad))
    assert rc == N.BT_E_SCHEMA
    with pytest.raises(RuntimeError, match="equal size"):
        sch.ingest(ring.ptr(slot), ring.view(slot), 0, bad)
    # CPU ring cannot submit
    with pytest.raises(N.NativeError):
        ring.submit(slot, 1234, 16, None)
    ring.close()


def test_schema_compat_path_for_strings():
    B = 3
    msgs = [{"btid": 0, "image": np.full((2, 4, 4), i, np.uint8), "label": "cube%d" % i} for i in range(B)]
    blobs = [pickle.dumps(m, protocol=3) for m in msgs]
    sch = Schema.from_message(blobs[0], B)
    assert not sch.native and sch.host_keys == ("label",)
    ring = Ring(sch.slot_bytes, depth=1, device=None)
    s = ring.acquire()
    rows = {}
    for i, b in enumerate(blobs):
        sch.ingest(ring.ptr(s), ring.view(s), i, b, rows)
    assert rows == {"label": ["cube0", "cube1", "cube2"]}
    c = sch.column("image")
    got = ring.view(s)[c.slot_off:c.slot_off + B * c.item_bytes].reshape(B, 2, 4, 4)
    np.testing.assert_array_equal(got, np.stack([m["image"] for m in msgs]))


def test_gamma_lut_matches_oracle(golden_dir):
    from oracle import preprocess
    tabs = np.load(golden_dir / "gamma_tables.npz")
    for name in tabs.files:
        g = float(name[1:].replace("_", "."))
        lut = np.frombuffer(N.gamma_lut(g), dtype=np.uint8)
        np.testing.assert_array_equal(lut, tabs[name])
    for g in (1.1, 1.5, 2.6, 3.0, 0.7):
        np.testing.assert_array_equal(np.frombuffer(N.gamma_lut(g), np.uint8), preprocess.gamma_table(g))


def test_shapes_whose_size_overflows_are_rejected_not_wrapped():
    """a crafted shape such as (2**32, 2**32, 3) multiplies to a small number modulo 2**64; it
    must not pass the payload-length comparison and become a schema column."""
    import pickle
    import numpy as np
    from blendtorch.btt import native as N
    from blendtorch.btt.staging import Schema
    good = pickle.dumps({"btid": 0, "image": np.zeros((2, 2, 3), np.uint8)}, protocol=3)
    def ints(dims):                                          # the opcodes that push the integers of a shape
        return b"".join(pickle.dumps(d, protocol=3)[2:-1] for d in dims)
    shape = ints((2, 2, 3)) + b"\x87"                          # ... followed by TUPLE3
    assert good.count(shape) == 1
    for dims in ((2 ** 32, 2 ** 32, 3), (2 ** 63 - 1, 2, 6), (2 ** 62, 4, 3)):
        bad = good.replace(shape, ints(dims) + b"\x87")
        f = {x.key.decode(): x for x in N.scan_message(bad)}
        assert f["image"].kind == N.BT_KIND_OTHER            # not an array the native path will touch
        h = __import__("ctypes").c_void_p()
        addr, n = N.buffer_address(bad)
        assert N.lib.bt_schema_create(addr, n, 8, __import__("ctypes").byref(h)) == N.BT_E_UNSUPPORTED
    # a marker (no payload to compare with) whose shape overflows cannot become a column
    huge = pickle.dumps({"btid": 0, "image": ("__btraw__", 1, "|u1", (2 ** 40, 2 ** 40))}, protocol=3)
    h = __import__("ctypes").c_void_p()
    addr, n = N.buffer_address(huge)
    assert N.lib.bt_schema_create(addr, n, 8, __import__("ctypes").byref(h)) == N.BT_E_TOO_LARGE
