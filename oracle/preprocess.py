"""Oracle: per-frame preprocessing and collation (TEST INFRASTRUCTURE).

The reference has no library function for preprocessing; the maths lives in the
producer's numpy tail and in the example `item_transform`s.  Those lines are the
specification of the fused CUDA kernel K1:

  flip       pkg_blender/blendtorch/btb/offscreen.py:95-96       np.flipud
  gamma      pkg_blender/blendtorch/btb/offscreen.py:105-112  ==  examples/datagen/generate.py:10-14
  normalise  examples/densityopt/densityopt.py:117-119           astype(f32); (x-127.5)/127.5; HWC->CHW
  collate    torch.utils.data.default_collate (collate.py:246-313)

Order when composed (hard part 2): flip -> gamma on uint8
-> truncate to uint8 -> float32 -> normalise -> transpose.  The uint8
truncation between gamma and normalise is part of the result and is kept.
"""

import numpy as np


def flip_vertical(frame):
    """offscreen.py:95-96: row y <-> H-1-y (materialised)."""
    return np.ascontiguousarray(frame[::-1])


def gamma_correct(frame, coeff=2.2):
    """offscreen.py:105-112 / generate.py:10-14.

    float32 arithmetic throughout (the array is float32 and the exponent is a
    Python scalar, so numpy keeps float32); `np.uint8(...)` truncates toward
    zero; only the first three channels are touched, alpha is appended back
    unchanged when present."""
    colour = frame[..., :3].astype(np.float32) / 255
    colour = np.uint8(255.0 * colour ** (1 / coeff))
    if frame.shape[-1] == 4:
        return np.concatenate((colour, frame[..., 3:4]), axis=-1)
    return colour


def gamma_table(coeff=2.2):
    """The 256 possible outputs of `gamma_correct` for one colour byte."""
    ramp = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(3, axis=2)
    return gamma_correct(ramp, coeff)[0, :, 0].copy()


def normalise_chw(frame, mean=127.5, std=127.5):
    """densityopt.py:117-119: float32 subtract then true divide, then HWC->CHW.

    `mean`/`std` generalise the literal 127.5 to per-channel float32 values
    (scalars or length-C sequences); with the defaults this is the reference
    expression exactly."""
    x = frame.astype(np.float32)
    m = np.asarray(mean, dtype=np.float32)
    s = np.asarray(std, dtype=np.float32)
    x = (x - m) / s
    return np.ascontiguousarray(np.transpose(x, (2, 0, 1)))


def preprocess_frame(frame, flip=False, gamma=None, mean=None, std=None,
                     chw=True, out_dtype="float32"):
    """Composition used by the parity tests; mirrors what a user would write as
    `item_transform` from the cited lines.  Returns a contiguous ndarray.

    out_dtype "uint8": flip/gamma only (no normalise), layout per `chw`.
    out_dtype "float32": astype(f32) then optional normalise.
    """
    x = frame
    if x.ndim == 2:
        x = x[..., None]
    if flip:
        x = flip_vertical(x)
    if gamma:
        x = gamma_correct(x, gamma)
    if out_dtype == "uint8":
        assert mean is None and std is None
        y = x
    else:
        y = x.astype(np.float32)
        if mean is not None or std is not None:
            m = np.asarray(0.0 if mean is None else mean, dtype=np.float32)
            s = np.asarray(1.0 if std is None else std, dtype=np.float32)
            y = (y - m) / s
    if chw:
        y = np.transpose(y, (2, 0, 1))
    return np.ascontiguousarray(y)


def preprocess_batch(frames, **kw):
    """Stack of `preprocess_frame` == what default_collate builds from the
    transformed items (collate.py:246-289: as_tensor + stack)."""
    return np.stack([preprocess_frame(f, **kw) for f in frames])


def batch_stats(batch_nchw):
    """Per-channel sum and sum of squares of an (B,C,H,W) float batch, float64
    accumulation.  Not in the reference (the north_star's optional per-batch
    statistics); oracle = the obvious numpy reduction."""
    x = batch_nchw.astype(np.float64)
    return x.sum(axis=(0, 2, 3)), (x * x).sum(axis=(0, 2, 3))


def to_bf16_bits(x):
    """float32 -> bfloat16 bit pattern, round-to-nearest-even (what
    torch.Tensor.to(torch.bfloat16) does); used for the 1-ulp bf16 checks."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    rounding = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + rounding) >> 16).astype(np.uint16)


def collate(items):
    """default_collate over a list of item dicts (third-party, by call)."""
    from torch.utils.data import default_collate

    return default_collate(items)
