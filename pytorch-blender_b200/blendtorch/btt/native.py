"""ctypes binding of libbt_b200.so (declared in include/bt_b200.h).

The library is the product: it is loaded from the tree (``_native/libbt_b200.so``,
built by ``build_native.py``) and there is deliberately no Python or CPU stand-in
for it.  If it is missing, importing this module raises.
"""
import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "_native" / "libbt_b200.so"

BT_OK = 0
BT_E_INVALID, BT_E_UNSUPPORTED, BT_E_TRUNCATED, BT_E_SCHEMA = -1, -2, -3, -4
BT_E_CUDA, BT_E_NOMEM, BT_E_IO, BT_E_TIMEOUT, BT_E_TOO_LARGE = -5, -6, -7, -8, -9

(BT_U8, BT_I8, BT_U16, BT_I16, BT_U32, BT_I32, BT_U64, BT_I64,
 BT_F16, BT_F32, BT_F64, BT_BOOL, BT_BF16) = range(13)
(BT_KIND_NDARRAY, BT_KIND_INT, BT_KIND_FLOAT, BT_KIND_BOOL,
 BT_KIND_NONE, BT_KIND_STR, BT_KIND_OTHER, BT_KIND_SHMREF, BT_KIND_RAWREF) = range(9)

BT_MAX_DIMS = 8
BT_MAX_KEY = 64

NUMPY_DTYPES = {
    BT_U8: "uint8", BT_I8: "int8", BT_U16: "uint16", BT_I16: "int16", BT_U32: "uint32",
    BT_I32: "int32", BT_U64: "uint64", BT_I64: "int64", BT_F16: "float16", BT_F32: "float32",
    BT_F64: "float64", BT_BOOL: "bool",
}
DTYPE_SIZE = {BT_U8: 1, BT_I8: 1, BT_U16: 2, BT_I16: 2, BT_U32: 4, BT_I32: 4, BT_U64: 8,
              BT_I64: 8, BT_F16: 2, BT_F32: 4, BT_F64: 8, BT_BOOL: 1, BT_BF16: 2}


class NativeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libbt_b200: %s (status %d)" % (msg, code))
        self.code = code


class Field(C.Structure):
    _fields_ = [("key", C.c_char * BT_MAX_KEY), ("kind", C.c_int32), ("dtype", C.c_int32),
                ("ndim", C.c_int32), ("fortran", C.c_int32), ("shape", C.c_int64 * BT_MAX_DIMS),
                ("payload_off", C.c_uint64), ("payload_len", C.c_uint64),
                ("ival", C.c_int64), ("fval", C.c_double)]


class Column(C.Structure):
    _fields_ = [("key", C.c_char * BT_MAX_KEY), ("kind", C.c_int32), ("dtype", C.c_int32),
                ("ndim", C.c_int32), ("reserved", C.c_int32), ("shape", C.c_int64 * BT_MAX_DIMS),
                ("item_bytes", C.c_uint64), ("slot_off", C.c_uint64)]


class PreprocessParams(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p),
                ("batch", C.c_int32), ("height", C.c_int32), ("width", C.c_int32), ("channels", C.c_int32),
                ("out_channels", C.c_int32), ("out_dtype", C.c_int32), ("flip", C.c_int32), ("chw", C.c_int32),
                ("use_gamma", C.c_int32), ("use_norm", C.c_int32), ("variant", C.c_int32), ("reserved", C.c_int32),
                ("src_image_stride", C.c_uint64),
                ("mean", C.c_float * 4), ("stdev", C.c_float * 4), ("gamma_lut", C.c_uint8 * 256),
                ("stats_out", C.c_void_p), ("stats_scratch", C.c_void_p),
                ("src_index", C.c_void_p)]


class ProjectParams(C.Structure):
    _fields_ = [("points", C.c_void_p), ("view", C.c_void_p), ("proj", C.c_void_p),
                ("batch", C.c_int32), ("npoints", C.c_int32),
                ("view_per_frame", C.c_int32), ("proj_per_frame", C.c_int32),
                ("height", C.c_int32), ("width", C.c_int32),
                ("origin_upper_left", C.c_int32), ("with_depth", C.c_int32),
                ("ndc_out", C.c_void_p), ("xy_out", C.c_void_p), ("depth_out", C.c_void_p),
                ("corners_per_object", C.c_int32), ("input_is_ndc", C.c_int32),
                ("bbox_out", C.c_void_p)]


class PushJob(C.Structure):
    _fields_ = [("first_endpoint", C.c_int32), ("n_endpoints", C.c_int32), ("pool_n", C.c_int32), ("reserved", C.c_int32),
                ("messages", C.c_int64), ("pool", C.c_void_p), ("pool_len", C.c_void_p),
                ("seg", C.c_void_p), ("payload", C.c_void_p), ("payload_len", C.c_void_p),
                ("tmpl_ticket_off", C.c_void_p), ("stop", C.c_void_p), ("sent", C.c_void_p)]


def _load():
    if not _LIB_PATH.exists():
        raise ImportError(
            "%s is missing: build it with `python pytorch-blender_b200/build_native.py` "
            "(there is no fallback implementation)" % _LIB_PATH)
    lib = C.CDLL(str(_LIB_PATH), mode=getattr(os, "RTLD_NOW", 2))
    vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
    sig = {
        "bt_last_error": (C.c_char_p, []),
        "bt_abi_version": (i32, []),
        "bt_scan_message": (i32, [vp, u64, C.POINTER(Field), i32]),
        "bt_schema_create": (i32, [vp, u64, i32, C.POINTER(vp)]),
        "bt_schema_destroy": (i32, [vp]),
        "bt_schema_num_columns": (i32, [vp]),
        "bt_schema_column": (i32, [vp, i32, C.POINTER(Column)]),
        "bt_schema_slot_bytes": (u64, [vp]),
        "bt_schema_ingest": (i32, [vp, vp, i32, vp, u64]),
        "bt_schema_ingest_many": (i32, [vp, vp, i32, i32, vp, vp, i32, C.POINTER(i32)]),
        "bt_ring_create": (i32, [i32, u64, i32, C.POINTER(vp)]),
        "bt_ring_destroy": (i32, [vp]),
        "bt_ring_depth": (i32, [vp]),
        "bt_ring_slot_bytes": (u64, [vp]),
        "bt_ring_slot_ptr": (vp, [vp, i32]),
        "bt_ring_acquire": (i32, [vp, C.POINTER(i32)]),
        "bt_ring_submit": (i32, [vp, i32, vp, u64, vp]),
        "bt_ring_wait": (i32, [vp, i32]),
        "bt_ring_ingest": (i32, [vp, i32, vp, i32, vp, u64]),
        "bt_ring_submit_batch": (i32, [vp, i32, vp, i32, vp, vp]),
        "bt_ring_spill": (i32, [vp, i32, vp]),
        "bt_ring_reap": (i32, [vp]),
        "bt_gather_h2d": (i32, [vp, vp, u64, vp, i32, vp]),
        "bt_rx_create": (i32, [vp, vp, C.POINTER(vp)]),
        "bt_rx_destroy": (i32, [vp]),
        "bt_rx_recv": (i32, [vp, i32]),
        "bt_rx_pending": (i32, [vp, C.POINTER(vp), C.POINTER(u64)]),
        "bt_rx_drop": (i32, [vp]),
        "bt_rx_fill": (i32, [vp, vp, i32, vp, i32, i32, i32, C.POINTER(i32)]),
        "bt_rx_release_queued": (i32, [vp, C.POINTER(i32)]),
        "bt_pull_create": (i32, [C.POINTER(vp)]),
        "bt_pull_connect": (i32, [vp, C.c_char_p]),
        "bt_pull_destroy": (i32, [vp]),
        "bt_pull_recv": (i32, [vp, i32]),
        "bt_pull_pending": (i32, [vp, C.POINTER(vp), C.POINTER(u64)]),
        "bt_pull_pending_parts": (i32, [vp, C.POINTER(u64), i32]),
        "bt_pull_drop": (i32, [vp]),
        "bt_pull_fill": (i32, [vp, vp, i32, vp, i32, i32, i32, C.POINTER(i32)]),
        "bt_push_create": (i32, [C.POINTER(vp)]),
        "bt_push_bind": (i32, [vp, C.c_char_p]),
        "bt_push_endpoint": (i32, [vp, i32, C.c_char_p, i32]),
        "bt_push_destroy": (i32, [vp]),
        "bt_push_peers": (i32, [vp, i32]),
        "bt_push_send": (i32, [vp, i32, vp, vp, i32, i32]),
        "bt_push_run": (i32, [vp, C.POINTER(PushJob)]),
        "bt_shm_create": (i32, [C.c_char_p, u64, i32, C.POINTER(vp)]),
        "bt_shm_attach": (i32, [C.c_char_p, i32, C.POINTER(vp)]),
        "bt_shm_close": (i32, [vp]),
        "bt_shm_nslots": (i32, [vp]),
        "bt_shm_slot_bytes": (u64, [vp]),
        "bt_shm_slot_ptr": (vp, [vp, i32]),
        "bt_shm_acquire": (i32, [vp, i32, C.POINTER(i32)]),
        "bt_shm_release": (i32, [vp, i32]),
        "bt_shm_check": (i32, [vp, i32]),
        "bt_shm_set_lease": (i32, [vp, C.c_int64]),
        "bt_build_gamma_lut": (i32, [C.c_double, C.POINTER(C.c_uint8 * 256)]),
        "bt_preprocess_scratch_bytes": (u64, [C.POINTER(PreprocessParams)]),
        "bt_preprocess_launch": (i32, [C.POINTER(PreprocessParams), vp]),
        "bt_preprocess_last_variant": (i32, []),
        "bt_preprocess_last_kernels": (i32, []),
        "bt_flip_rows_launch": (i32, [vp, vp, i32, i32, u64, vp, vp]),
        "bt_preprocess_stats_static": (i32, [C.POINTER(PreprocessParams), vp]),
        "bt_project_launch": (i32, [C.POINTER(ProjectParams), vp]),
        "bt_device_count": (i32, [C.POINTER(i32)]),
        "bt_measure_h2d": (i32, [i32, u64, i32, C.POINTER(C.c_double)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()
LIB_PATH = _LIB_PATH
EXPORTS = ["bt_last_error", "bt_abi_version", "bt_scan_message", "bt_schema_create", "bt_schema_destroy",
           "bt_schema_num_columns", "bt_schema_column", "bt_schema_slot_bytes", "bt_schema_ingest",
           "bt_schema_ingest_many", "bt_ring_create", "bt_ring_destroy", "bt_ring_depth",
           "bt_ring_slot_bytes", "bt_ring_slot_ptr", "bt_ring_acquire", "bt_ring_submit", "bt_ring_wait",
           "bt_ring_ingest", "bt_ring_submit_batch", "bt_ring_spill", "bt_ring_reap", "bt_gather_h2d",
           "bt_rx_create", "bt_rx_destroy", "bt_rx_recv", "bt_rx_pending", "bt_rx_drop", "bt_rx_fill",
           "bt_rx_release_queued",
           "bt_pull_create", "bt_pull_connect", "bt_pull_destroy", "bt_pull_recv", "bt_pull_pending", "bt_pull_pending_parts",
           "bt_pull_drop", "bt_pull_fill",
           "bt_push_create", "bt_push_bind", "bt_push_endpoint", "bt_push_destroy", "bt_push_peers", "bt_push_send",
           "bt_push_run",
           "bt_shm_create", "bt_shm_attach", "bt_shm_close", "bt_shm_nslots", "bt_shm_slot_bytes",
           "bt_shm_slot_ptr", "bt_shm_acquire", "bt_shm_release", "bt_shm_check", "bt_shm_set_lease",
           "bt_build_gamma_lut", "bt_preprocess_scratch_bytes", "bt_preprocess_launch",
           "bt_preprocess_last_variant", "bt_preprocess_last_kernels", "bt_flip_rows_launch", "bt_preprocess_stats_static", "bt_project_launch", "bt_raster_launch", "bt_device_count", "bt_measure_h2d"]


def last_error():
    return (lib.bt_last_error() or b"").decode("utf-8", "replace")


def check(rc):
    if rc < 0:
        raise NativeError(rc, last_error())
    return rc


def buffer_address(buf):
    """(address, nbytes) of any object exporting the buffer protocol, without copying."""
    mv = memoryview(buf)
    if not mv.contiguous:
        raise ValueError("buffer must be contiguous")
    n = mv.nbytes
    if n == 0:
        return 0, 0
    if mv.readonly:
        # ctypes.from_buffer needs a writable view; numpy reads the address of read-only ones
        import numpy as np
        arr = np.frombuffer(mv, dtype=np.uint8)
        return arr.ctypes.data, n
    return C.addressof(C.c_char.from_buffer(mv)), n


def scan_message(buf, max_fields=64):
    """List of Field structs for a pickled message; raises NativeError(BT_E_UNSUPPORTED)
    when the pickle is outside the ingest subset."""
    addr, n = buffer_address(buf)
    out = (Field * max_fields)()
    cnt = check(lib.bt_scan_message(addr, n, out, max_fields))
    return [out[i] for i in range(cnt)]


def gamma_lut(coeff):
    out = (C.c_uint8 * 256)()
    check(lib.bt_build_gamma_lut(float(coeff), C.byref(out)))
    return bytes(out)
