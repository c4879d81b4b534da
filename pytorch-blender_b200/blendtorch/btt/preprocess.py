"""Device-side frame preprocessing: the host face of kernel K1.

What users of the reference write as a numpy ``item_transform`` --
gamma (examples/datagen/generate.py:10-14 == btb/offscreen.py:105-112), normalise + HWC->CHW
(examples/densityopt/densityopt.py:117-119), flip (btb/offscreen.py:95-96) -- is described
here declaratively by `Preprocess` and executed for a whole batch by one launch of
``bt_preprocess_launch`` on the GPU.  There is no CPU implementation in this package.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence, Union

import torch

from . import native as N

_TORCH_OUT = {torch.uint8: N.BT_U8, torch.float32: N.BT_F32, torch.bfloat16: N.BT_BF16}
VARIANTS = {"auto": 0, "generic": 1, "vector": 2, "bulk": 3, "pixel": 4}


def _per_channel(v, default):
    if v is None:
        return [default] * 4
    if isinstance(v, (int, float)):
        return [float(v)] * 4
    v = [float(x) for x in v]
    if not 1 <= len(v) <= 4:
        raise ValueError("per-channel values must have 1..4 entries")
    return v + [default] * (4 - len(v))


@dataclass
class Preprocess:
    """What to do with the uint8 frames of one message field.

    key:        dict key of the image field (``'image'`` in the reference's examples).
    flip:       vertical flip (``np.flipud``): for frames sent bottom-up (origin='lower-left').
    gamma:      gamma coefficient, e.g. 2.2; applied to the first three channels only and
                truncated to uint8 exactly like the reference; None = off.
    mean, std:  per-channel (or scalar) normalisation ``(x - mean) / std`` in float32;
                both None = off (plain cast).
    layout:     'NCHW' (the reference's ``np.transpose(x, (2, 0, 1))``) or 'NHWC'.
    dtype:      torch.float32, torch.bfloat16 or torch.uint8 (flip/gamma only).
    out_channels: keep only the first k channels (e.g. 3 drops alpha); None = all.
    stats:      also return per-channel [sum, sum of squares] of the batch (float64).
    """
    key: str = "image"
    flip: bool = False
    gamma: Optional[float] = None
    mean: Union[None, float, Sequence[float]] = None
    std: Union[None, float, Sequence[float]] = None
    layout: str = "NCHW"
    dtype: torch.dtype = torch.float32
    out_channels: Optional[int] = None
    stats: bool = False
    variant: str = "auto"
    tune: int = 0   # developer knob: low byte = units per thread, next byte = CTAs per SM cap (0 = defaults)
    _lut: bytes = field(default=b"", repr=False)

    def __post_init__(self):
        if self.layout not in ("NCHW", "NHWC"):
            raise ValueError("layout must be 'NCHW' or 'NHWC'")
        if self.dtype not in _TORCH_OUT:
            raise ValueError("dtype must be torch.float32, torch.bfloat16 or torch.uint8")
        if self.dtype == torch.uint8 and (self.mean is not None or self.std is not None):
            raise ValueError("normalisation needs a floating output dtype")
        if self.variant not in VARIANTS:
            raise ValueError("variant must be one of %s" % sorted(VARIANTS))
        if self.gamma:
            self._lut = N.gamma_lut(self.gamma)


def _stream_handle(stream):
    if stream is None:
        stream = torch.cuda.current_stream()
    return C.c_void_p(stream.cuda_stream)


def preprocess_images(src, spec, out=None, stream=None, stats_out=None, _scratch=None, index=None):
    """Run K1 on `src`, a CUDA uint8 tensor (B,H,W,C) or (B,H,W), per `spec`.

    index: optional CUDA int64 tensor (n,): output item i is made from ``src[index[i]]`` -- a batch
        drawn from frames resident on the device, without a gather pass.

    Returns the output tensor, or (output, stats) when ``spec.stats``; stats is a
    (C_out, 2) float64 CUDA tensor [sum, sum of squares].  Work is enqueued on `stream`
    (default: torch's current stream); nothing is synchronised.
    """
    if not src.is_cuda:
        raise RuntimeError("preprocess_images needs a CUDA tensor; this package has no CPU kernels")
    if src.dtype != torch.uint8:
        raise TypeError("frames must be uint8, got %s" % src.dtype)
    if src.dim() == 3:
        src = src.unsqueeze(-1)
    if src.dim() != 4:
        raise ValueError("expected (B,H,W,C) frames")
    B, H, W, Cn = src.shape
    if not 1 <= Cn <= 4:
        raise ValueError("channels must be 1..4")
    if src.stride(3) != 1 or src.stride(2) != Cn or src.stride(1) != W * Cn:
        src = src.contiguous()
    n_src = B
    if index is not None:
        if not index.is_cuda or index.dtype != torch.int64 or index.dim() != 1 or not index.is_contiguous():
            raise ValueError("index must be a contiguous 1-d CUDA int64 tensor")
        if B == 0 and index.numel():
            raise IndexError("index into an empty set of frames")
        B = index.numel()
    cout = spec.out_channels or Cn
    if not 1 <= cout <= Cn:
        raise ValueError("out_channels must be within 1..channels")
    shape = (B, cout, H, W) if spec.layout == "NCHW" else (B, H, W, cout)
    if out is None:
        out = torch.empty(shape, dtype=spec.dtype, device=src.device)
    else:
        if tuple(out.shape) != shape or out.dtype != spec.dtype or not out.is_contiguous():
            raise ValueError("out must be a contiguous %s tensor of shape %s" % (spec.dtype, shape))

    if out.numel() == 0:  # nothing to launch for an empty batch
        if spec.stats:
            st = stats_out if stats_out is not None else torch.empty((cout, 2), dtype=torch.float64, device=src.device)
            return out, st.zero_()
        return out

    p = N.PreprocessParams()
    p.src = src.data_ptr()
    p.dst = out.data_ptr()
    p.batch, p.height, p.width, p.channels = B, H, W, Cn
    p.out_channels = cout
    p.out_dtype = _TORCH_OUT[spec.dtype]
    p.flip = int(bool(spec.flip))
    p.chw = int(spec.layout == "NCHW")
    p.use_gamma = int(bool(spec.gamma))
    p.use_norm = int(spec.mean is not None or spec.std is not None)
    p.variant = VARIANTS[spec.variant]
    p.reserved = int(spec.tune)
    p.src_image_stride = src.stride(0) if n_src > 1 else 0
    p.src_index = index.data_ptr() if index is not None else None
    mean, std = _per_channel(spec.mean, 0.0), _per_channel(spec.std, 1.0)
    for c in range(4):
        p.mean[c], p.stdev[c] = mean[c], std[c]
    if spec.gamma:
        C.memmove(p.gamma_lut, spec._lut, 256)
    stats = None
    if spec.stats:
        stats = stats_out if stats_out is not None else torch.empty((cout, 2), dtype=torch.float64, device=src.device)
        p.stats_out = stats.data_ptr()
        nbytes = N.lib.bt_preprocess_scratch_bytes(C.byref(p))
        scratch = _scratch
        if scratch is None or scratch.numel() < nbytes:
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=src.device)
        p.stats_scratch = scratch.data_ptr()
        if stream is not None:
            scratch.record_stream(stream)
    with torch.cuda.device(src.device):
        N.check(N.lib.bt_preprocess_launch(C.byref(p), _stream_handle(stream)))
    return (out, stats) if spec.stats else out


class PreparedPreprocess:
    """`preprocess_images` for a fixed frame shape with everything that does not change from batch
    to batch (parameter struct, gamma table, statistics scratch) set up once -- the per-batch cost
    of the host side is two allocations and one ctypes call.  Used by DeviceLoader."""

    def __init__(self, spec, height, width, channels, device):
        if not 1 <= channels <= 4:
            raise ValueError("channels must be 1..4")
        self.spec, self.device = spec, device
        self.h, self.w, self.c = height, width, channels
        self.cout = spec.out_channels or channels
        if not 1 <= self.cout <= channels:
            raise ValueError("out_channels must be within 1..channels")
        p = N.PreprocessParams()
        p.height, p.width, p.channels, p.out_channels = height, width, channels, self.cout
        p.out_dtype = _TORCH_OUT[spec.dtype]
        p.flip, p.chw = int(bool(spec.flip)), int(spec.layout == "NCHW")
        p.use_gamma = int(bool(spec.gamma))
        p.use_norm = int(spec.mean is not None or spec.std is not None)
        p.variant, p.reserved = VARIANTS[spec.variant], int(spec.tune)
        p.src_image_stride = 0
        mean, std = _per_channel(spec.mean, 0.0), _per_channel(spec.std, 1.0)
        for c in range(4):
            p.mean[c], p.stdev[c] = mean[c], std[c]
        if spec.gamma:
            C.memmove(p.gamma_lut, spec._lut, 256)
        self.params = p
        self._ref = C.byref(p)
        self.scratch = None
        if spec.stats:
            p.stats_out = 1  # non-null so that the size query answers for the statistics path
            self.scratch = torch.empty(N.lib.bt_preprocess_scratch_bytes(self._ref), dtype=torch.uint8, device=device)
            p.stats_scratch = self.scratch.data_ptr()
            p.stats_out = None

    def out_shape(self, n):
        return (n, self.cout, self.h, self.w) if self.spec.layout == "NCHW" else (n, self.h, self.w, self.cout)

    def run(self, src_ptr, n, stream_handle, index_ptr=None):
        """Launch on `n` dense frames at device address `src_ptr` (with `index_ptr`, the device
        address of `n` int64: on frames ``index[i]`` of the dense set at `src_ptr`); allocations come
        from torch's *current* stream, which the caller has set to the stream behind `stream_handle`."""
        out = torch.empty(self.out_shape(n), dtype=self.spec.dtype, device=self.device)
        stats = None
        p = self.params
        if self.spec.stats:
            stats = torch.empty((self.cout, 2), dtype=torch.float64, device=self.device)
            p.stats_out = stats.data_ptr()
        self.last_kernels = 0
        if n == 0:
            return out, (stats.zero_() if stats is not None else None)
        p.src, p.dst, p.batch = src_ptr, out.data_ptr(), n
        p.src_index = index_ptr
        N.check(N.lib.bt_preprocess_launch(self._ref, stream_handle))
        self.last_kernels = 2 if stats is not None else 1   # K1 + the fold of its statistics rows
        if stats is not None:
            self.last_kernels = N.lib.bt_preprocess_last_kernels()
        return out, stats


    def run_group(self, src_ptr, n, batch, stream_handle):
        """`n` dense frames that will be handed out as batches of `batch`: one output allocation for all
        of them and, without statistics, one launch; with statistics one launch per batch (they are
        per-batch results), writing into slices of the same output.  Returns (out, stats) with stats of
        shape (number of batches, C_out, 2) or None."""
        out = torch.empty(self.out_shape(n), dtype=self.spec.dtype, device=self.device)
        self.last_kernels = 0
        p = self.params
        p.src_index = None
        if not self.spec.stats:
            if n:
                p.src, p.dst, p.batch = src_ptr, out.data_ptr(), n
                N.check(N.lib.bt_preprocess_launch(self._ref, stream_handle))
                self.last_kernels = 1
            return out, None
        nb = max(1, -(-n // batch))
        stats = torch.empty((nb, self.cout, 2), dtype=torch.float64, device=self.device)
        if n == 0:
            return out, stats.zero_()
        in_item = self.h * self.w * self.c
        out_item = self.h * self.w * self.cout * out.element_size()
        dst, st = out.data_ptr(), stats.data_ptr()
        for b in range(nb):
            first = b * batch
            p.src, p.dst = src_ptr + first * in_item, dst + first * out_item
            p.batch = min(batch, n - first)
            p.stats_out = st + b * self.cout * 16
            N.check(N.lib.bt_preprocess_launch(self._ref, stream_handle))
            self.last_kernels += N.lib.bt_preprocess_last_kernels()
        return out, stats


@dataclass
class FlipRows:
    """Vertical flip of a field K1 does not take -- a float32 depth map, a uint16 mask -- sent bottom-up.
    The colour buffer of the reference is flipped on the producer (btb/offscreen.py:95-96); a depth
    read-back has the same order.  One copy at HBM speed (``bt_flip_rows_launch``), any dtype and shape
    (items are (H, ...) arrays, rows are everything after the first axis)."""
    key: str = "depth"


def flip_rows(src, out=None, stream=None, index=None):
    """``torch.flip(src, dims=[1])`` for a contiguous CUDA tensor (B, H, ...) -- or, with `index`, of
    ``src[index]`` -- by the library's kernel."""
    if not src.is_cuda or not src.is_contiguous() or src.dim() < 2:
        raise ValueError("flip_rows takes a contiguous CUDA tensor of shape (B, H, ...)")
    n = src.shape[0] if index is None else index.numel()
    if index is not None and (not index.is_cuda or index.dtype != torch.int64 or not index.is_contiguous()):
        raise ValueError("index must be a contiguous CUDA int64 tensor")
    shape = (n,) + tuple(src.shape[1:])
    if out is None:
        out = torch.empty(shape, dtype=src.dtype, device=src.device)
    elif tuple(out.shape) != shape or out.dtype != src.dtype or not out.is_contiguous():
        raise ValueError("out must be a contiguous %s tensor of shape %s" % (src.dtype, shape))
    if out.numel():
        h = src.shape[1]
        row_bytes = src[0, 0].numel() * src.element_size() if src.dim() > 2 else src.element_size()
        with torch.cuda.device(src.device):
            N.check(N.lib.bt_flip_rows_launch(src.data_ptr(), out.data_ptr(), n, h, row_bytes,
                                              index.data_ptr() if index is not None else None, _stream_handle(stream)))
    return out


class PreparedFlip:
    """`flip_rows` for a fixed item shape; same calling convention as `PreparedPreprocess`."""

    def __init__(self, spec, shape, dtype, device):
        if len(shape) < 1:
            raise ValueError("field %r is a scalar; there are no rows to flip" % spec.key)
        self.spec, self.device, self.shape, self.dtype = spec, device, tuple(shape), dtype
        self.h = shape[0]
        self.row_bytes = int(torch.empty(0, dtype=dtype).element_size())
        for d in shape[1:]:
            self.row_bytes *= d
        self.last_kernels = 0

    def run_group(self, src_ptr, n, batch, stream_handle):
        return self.run(src_ptr, n, stream_handle)

    def run(self, src_ptr, n, stream_handle, index_ptr=None):
        out = torch.empty((n,) + self.shape, dtype=self.dtype, device=self.device)
        self.last_kernels = 0
        if out.numel():
            N.check(N.lib.bt_flip_rows_launch(src_ptr, out.data_ptr(), n, self.h, self.row_bytes, index_ptr, stream_handle))
            self.last_kernels = 1
        return out, None


def prepare(spec, shape, np_dtype, torch_dtype, device):
    """The per-stream launcher of a field's spec (`Preprocess` or `FlipRows`) for items of `shape`."""
    import numpy as np
    if isinstance(spec, FlipRows):
        return PreparedFlip(spec, shape, torch_dtype, device)
    if np_dtype != np.uint8:
        raise TypeError("field %r is %s; Preprocess takes uint8 frames (FlipRows flips anything)" % (spec.key, np_dtype))
    shp = tuple(shape) if len(shape) == 3 else tuple(shape) + (1,)
    if len(shp) != 3:
        raise ValueError("field %r has shape %s; frames are (H,W,C) or (H,W)" % (spec.key, list(shape)))
    return PreparedPreprocess(spec, shp[0], shp[1], shp[2], device)
