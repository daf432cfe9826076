"""Name-compatible stand-in for the reference's native module (M/COALA_GNN_Pybind.cu:27-79): the ctypes binding of
libcoala_b200.so by default, the compiled pybind11 module (COALA_GNN_Pybind/_native*.so, built from
coala_gnn_b200/csrc_pybind/COALA_GNN_Pybind.cpp) with COALA_PYBIND_NATIVE=1."""
import os

if os.environ.get("COALA_PYBIND_NATIVE", "0") == "1":
    from coala_gnn_b200 import binding as _binding  # builds the module if needed
    from ._native import *  # noqa: F401,F403
    from coala_gnn_b200.pybind import tensor_from_ptr  # noqa: F401
else:
    from coala_gnn_b200.pybind import *  # noqa: F401,F403
    from coala_gnn_b200.pybind import __all__  # noqa: F401
