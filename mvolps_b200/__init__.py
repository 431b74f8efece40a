"""mvolps_b200 -- B200-native LP-relaxation engine behind MVOLPS' GLPK boundary.

Python is only the test/bench harness around the C ABIs (include/*.h); the
product is mvolps_b200/lib/libglpk_b200.so (C++ host + sm_100a CUDA kernels).
"""
from . import capi  # noqa: F401
from .lp import EngineLP, GlpProb, solve_batch  # noqa: F401
from . import instances  # noqa: F401
