"""`import pypolar` -- the reference's module name, served by the B200 mirror (polar-codes_b200/pypolar.py over
libpolar_b200.so). Put the repository root on sys.path and existing scripts written for the reference's pybind module
(python/__init__.py:16-21) run unchanged: PolarEncoder, PolarDecoder, Detector, Puncturer, frozen_bits."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import polar_b200_loader  # noqa: E402

_pkg = polar_b200_loader.load()
from polar_codes_b200.pypolar import *  # noqa: E402,F401,F403
from polar_codes_b200.pypolar import Detector, PolarDecoder, PolarEncoder, Puncturer, frozen_bits  # noqa: E402,F401
