"""htr-solver_b200: B200-native right-hand-side engine for the HTR solver (one hot path).
Import as `htr_solver_b200` (the hyphenated directory is the package body)."""
from .core import *  # noqa: F401,F403
from .core import Solver, Fluid, HTRError, load_library, lib_path, SCHEMES, FIELD_SPECS, Rect
from . import synthetic
