"""B200-native batched polar codec: hot path of ant-uni-bremen/polar-codes as sm_100a CUDA kernels behind a
C-ABI (include/polar_b200.h) and a pypolar-compatible host interface (pypolar.py).

The directory name carries a hyphen (it follows the reference's repository name), so import it through
`polar_b200_loader.load()` at the repository root, or add it to sys.modules under `polar_codes_b200`.
"""
from . import capi, pcsim, pypolar, sweep  # noqa: F401
from .pypolar import Detector, PolarDecoder, PolarEncoder, Puncturer, frozen_bits  # noqa: F401
