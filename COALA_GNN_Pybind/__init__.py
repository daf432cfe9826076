"""Name-compatible stand-in for the reference's native module (M/COALA_GNN_Pybind.cu:27-79)."""
from coala_gnn_b200.pybind import *  # noqa: F401,F403
from coala_gnn_b200.pybind import __all__  # noqa: F401
