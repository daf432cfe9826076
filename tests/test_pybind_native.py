"""The compiled pybind11 binding (COALA_GNN_Pybind/_native*.so, csrc_pybind/COALA_GNN_Pybind.cpp) exports the reference
module's interface (M/COALA_GNN_Pybind.cu:27-79): same seven classes, same method names.  CPU-only: no compute call."""
import pytest

REFERENCE_API = {  # class -> methods bound at M/COALA_GNN_Pybind.cu:27-79
    "SharedUVAManager": ["get_host_ptr", "get_device_ptr", "cleanup"],
    "SSD_GNN_SSD_Controllers": [],
    "SSD_GNN_NVSHMEM_Cache": ["send_requests", "read_feature", "get_cache_data", "print_stats"],
    "Isolated_Cache": ["read_feature", "get_cache_data", "split_node_list", "nccl_get_feature", "map_feat_data", "print_stats"],
    "Node_distributor_pybind": ["distribute_node_with_affinity", "get_num_colors"],
    "NVSHMEM_Manager": ["allocate", "free", "finalize"],
    "Graph_Coloring": ["cpu_color_graph", "cpu_color_graph_optimized", "cpu_count_nearest_color", "cpu_count_nearest_color_less_memory",
                       "cpu_calculate_color_affinity", "set_color_buffer", "set_topk_color_buffer", "set_topk_affinity_buffer", "set_adj_csc",
                       "get_num_color_node", "get_num_color"],
}


@pytest.fixture(scope="module")
def native():
    from coala_gnn_b200 import build
    build.build_pybind()
    import COALA_GNN_Pybind._native as m
    return m


def test_native_module_exports_reference_classes_and_methods(native):
    for cls, methods in REFERENCE_API.items():
        assert hasattr(native, cls), cls
        for meth in methods:
            assert hasattr(getattr(native, cls), meth), "%s.%s" % (cls, meth)
    assert native.abi_version() == 1


def test_ctypes_binding_exports_the_same_interface():
    from coala_gnn_b200 import pybind
    for cls, methods in REFERENCE_API.items():
        for meth in methods:
            assert hasattr(getattr(pybind, cls), meth), "%s.%s" % (cls, meth)


def test_native_errors_are_runtime_errors(native):
    with pytest.raises(RuntimeError):
        native.SSD_GNN_SSD_Controllers(1, 4096, 1024, 0, 0, 2048, True)  # "Only Feature Embedding Size less than 8KB is supported"
    with pytest.raises(RuntimeError):
        native.SSD_GNN_SSD_Controllers(1, 4096, 1024, 0, 0, 1024, False)  # the BaM/NVMe tier is out of scope
    ctrl = native.SSD_GNN_SSD_Controllers(1, 4096, 1024, 0, 0, 100, True)
    assert (ctrl.cache_dim, ctrl.page_size) == (128, 512)  # M/ssd_gnn_cache.cuh:34-47
