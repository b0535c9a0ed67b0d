"""Importable alias of the package directory `htr-solver_b200/` (a hyphen cannot be imported)."""
import os as _os
__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "htr-solver_b200"))
from .core import *  # noqa: F401,F403,E402
from .core import Solver, Fluid, HTRError, load_library, lib_path, SCHEMES, FIELD_SPECS, Rect, mixture_bytes  # noqa: E402
from . import synthetic, build  # noqa: E402,F401
