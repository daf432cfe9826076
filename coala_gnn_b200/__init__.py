"""coala_gnn_b200 — B200-native (sm_100a) mini-batch data path of COALA-GNN:
colour-affinity seed distribution -> UVA neighbour sampling -> cached multi-GPU feature gather.
The work is done by libcoala_b200.so (C ABI in include/coala_b200.h); this package is the
host-side mirror of the reference's COALA_GNN_Pybind / COALA_GNN interfaces. No CPU fallback."""
__version__ = "0.1.0"
