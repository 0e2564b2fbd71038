"""doma_b200 — B200-native (sm_100a CUDA) S_N transport sweep hot path behind DOMA's driver API.

Only what the path needs lives here: ``csrc/`` (CUDA kernels + the C ABI of include/doma_b200.h),
``abi.py`` (ctypes binding), ``input_parameters.py`` / ``quadrature.py`` (mirror of the reference's
input API), ``solver.py`` (mirror of its outer loops), ``sharding.py`` (ordinate sharding + the
phi all-reduce) and ``driver/`` (the reference's own C++ loops bound to the ABI).
"""
from . import abi
from .abi import Context, DomaError, Problem, load_library
from .input_parameters import InputError, InputParameters, build_problem, load_case, parse_input_parameters
from .solver import TransportSolver, Trace

__all__ = ["abi", "Context", "DomaError", "Problem", "load_library", "InputError", "InputParameters",
           "build_problem", "load_case", "parse_input_parameters", "TransportSolver", "Trace"]
