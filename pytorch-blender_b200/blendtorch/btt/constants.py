"""Constants shared with the reference (pkg_pytorch/blendtorch/btt/constants.py)."""

# Default socket timeout: 10 s (constants.py:4).
DEFAULT_TIMEOUTMS = 10000
