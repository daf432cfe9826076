"""Name-compatible stand-in for the reference's Python package (S/__init__.py:1-3)."""
from coala_gnn_b200.comm import MPI_Comm_Manager, Shared_UVA_Tensor_Manager  # noqa: F401
from coala_gnn_b200.node_distributor import Node_Distributor  # noqa: F401
from coala_gnn_b200.dataloader import SSD_INFO, COALA_GNN_DataLoader, COALA_GNN_Node_Distribution_Scheduler  # noqa: F401
from coala_gnn_b200.manager import COALA_GNN_Manager  # noqa: F401
from coala_gnn_b200.sampler import MultiLayerNeighborSampler, UVAGraph, Block  # noqa: F401
