"""su2_b200 — B200-native implicit finite-volume flow iteration for SU2 (C ABI in include/su2b200.h)."""
from . import capi  # noqa: F401
from .solver import FlowSolver, LinearSystem, default_config  # noqa: F401
