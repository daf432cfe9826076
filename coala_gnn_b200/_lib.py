"""ctypes binding of libcoala_b200.so (the C ABI declared in include/coala_b200.h).

There is no CPU fallback: if the library is missing and cannot be built, or a call fails, this
raises.  ctypes releases the GIL for the duration of every call, which the reference's pybind11
module does not (SURVEY.md §8b "Threading").
"""
import ctypes as C
import os

from . import build as _build

_u64, _i64, _i32, _vp, _cp = C.c_uint64, C.c_int64, C.c_int, C.c_void_p, C.c_char_p
_dbl = C.c_double

# name: (restype, [argtypes])   — one entry per declaration in include/coala_b200.h
SIGNATURES = {
    "coala_last_error": (_cp, []),
    "coala_abi_version": (_i32, []),
    "coala_uva_create": (_vp, [_cp, _i64, _i32, _i32, _i32]),
    "coala_uva_host_ptr": (_u64, [_vp]),
    "coala_uva_device_ptr": (_u64, [_vp]),
    "coala_uva_cleanup": (_i32, [_vp, _i32]),
    "coala_uva_destroy": (None, [_vp]),
    "coala_nd_create": (_vp, [_u64, _i32, _i32, _i32, _i32, _cp, _cp, _cp, _i32]),
    "coala_nd_create_from_arrays": (_vp, [_u64, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _i32, _i32, _i32]),
    "coala_nd_create_plain": (_vp, [_u64, _i32]),
    "coala_nd_distribute": (_i32, [_vp, _u64, _u64, _vp, _i32]),
    "coala_nd_distribute_async": (_i32, [_vp, _u64, _u64, _vp, _i32, _vp]),
    "coala_nd_distribute_async_slot": (_i32, [_vp, _u64, _u64, _vp, _i32, _i32, _vp]),
    "coala_nd_set_meta_len": (_i32, [_vp, _i32]),
    "coala_nd_last_bucket_len": (_i32, [_vp, _vp]),
    "coala_nd_num_colors": (_i32, [_vp]),
    "coala_nd_color_host_ptr": (_u64, [_vp]),
    "coala_nd_num_graph_nodes": (_i64, [_vp]),
    "coala_nd_destroy": (None, [_vp]),
    "coala_cache_dim_for": (_i32, [_i32]),
    "coala_cache_create": (_vp, [_i32, _i32, _i32, _vp, _i32, _i32, _i32, _u64, _u64]),
    "coala_cache_destroy": (None, [_vp]),
    "coala_cache_num_sets": (_u64, [_vp]),
    "coala_cache_row_bytes": (_i32, [_vp]),
    "coala_cache_ipc_handle": (_i32, [_vp, _vp]),
    "coala_cache_attach_ipc": (_i32, [_vp, _i32, _vp]),
    "coala_cache_attach_local": (_i32, [_vp, _i32, _vp]),
    "coala_cache_is_fd_shared": (_i32, [_vp]),
    "coala_cache_export_fd": (_i32, [_vp]),
    "coala_cache_attach_fd": (_i32, [_vp, _i32, _i32]),
    "coala_cache_read_feature": (_i32, [_vp, _u64, _u64, _i64]),
    "coala_cache_read_feature_async": (_i32, [_vp, _u64, _u64, _i64, _u64, _vp]),
    "coala_cache_set_num_nodes": (_i32, [_vp, _i64]),
    "coala_cache_refused_ids": (_i32, [_vp, _vp, _i32]),
    "coala_cache_read_feature_serial": (_i32, [_vp, _u64, _u64, _i64]),
    "coala_cache_get_cache_data": (_i32, [_vp, _u64]),
    "coala_cache_get_cache_data_n": (_i32, [_vp, _u64, _i32]),
    "coala_cache_get_cache_data_async": (_i32, [_vp, _u64, _i32, _vp]),
    "coala_cache_color_counters_dev_ptr": (_u64, [_vp]),
    "coala_cache_stats": (_i32, [_vp, _vp, _vp, _vp, _i32]),
    "coala_cache_print_stats": (_i32, [_vp]),
    "coala_cache_export_state": (_i32, [_vp, _vp, _vp, _vp]),
    "coala_cache_set_copy_engine": (_i32, [_vp, _i32]),
    "coala_split_node_list": (_i32, [_u64, _i64, _u64, _u64, _u64, _i32, _i32, _vp]),
    "coala_split_node_list_static": (_i32, [_u64, _i64, _u64, _u64, _u64, _i32, _i32, _vp]),
    "coala_split_node_list_static_max_ids": (_i64, []),
    "coala_map_feat_data": (_i32, [_u64, _u64, _u64, _i64, _i32, _vp]),
    "coala_nccl_unique_id": (_i32, [_vp]),
    "coala_nccl_version": (_i32, []),
    "coala_nccl_create": (_vp, [_vp, _i32, _vp, _i32, _i32, _i64, _i32]),
    "coala_nccl_destroy": (None, [_vp]),
    "coala_nccl_fetch_async": (_i32, [_vp, _u64, _u64, _i64, _u64, _vp]),
    "coala_nccl_counters": (_i32, [_vp, _vp, _vp, _vp]),
    "coala_sampler_create": (_vp, [_i32, _u64, _u64, _u64, _i64, _i32, _vp, _i32]),
    "coala_sampler_destroy": (None, [_vp]),
    "coala_sampler_max_sample": (_i64, [_vp]),
    "coala_sampler_stage_indptr": (_i32, [_vp]),
    "coala_sampler_stage_indices": (_i32, [_vp]),
    "coala_sampler_share_indptr": (_i32, [_vp, _vp]),
    "coala_sampler_trace": (_i32, [_vp, _vp, _i32]),
    "coala_sampler_graph_nodes": (_i32, [_vp]),
    "coala_sampler_sample_async": (_i32, [_vp, _u64, _i32, _u64, _vp]),
    "coala_sampler_counts": (_i32, [_vp, _vp]),
    "coala_sampler_layer_ptrs": (_i32, [_vp, _i32, _vp, _vp, _vp, _vp]),
    "coala_sampler_input_count_dev_ptr": (_u64, [_vp]),
    "coala_lookup_edge_ids": (_i32, [_u64, _u64, _i64, _u64, _vp]),
    "coala_batch_key": (_u64, [_u64, _u64, _i32]),
    "coala_engine_create": (_vp, [_i32, _i32, _i32, _u64, _i32, _i32]),
    "coala_engine_destroy": (None, [_vp]),
    "coala_engine_bind_slot": (_i32, [_vp, _i32, _vp, _u64]),
    "coala_engine_bind_gather": (_i32, [_vp, _vp, _vp, _i64]),
    "coala_engine_set_seed": (_i32, [_vp, _u64]),
    "coala_engine_slot_stream": (_u64, [_vp, _i32]),
    "coala_engine_slot_seeds_dev": (_u64, [_vp, _i32]),
    "coala_engine_gather_stream": (_u64, [_vp]),
    "coala_engine_checkpoint": (_i32, [_vp, _vp]),
    "coala_engine_begin": (_i32, [_vp, _i64, _vp]),
    "coala_engine_sample_gather": (_i32, [_vp, _i64, _u64, _i32]),
    "coala_engine_launch_host": (_i32, [_vp, _i64, _u64, _i32, _vp]),
    "coala_engine_collect": (_i32, [_vp, _i64, _vp, _vp]),
    "coala_engine_slot_of": (_i32, [_vp, _i64]),
    "coala_engine_times": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32]),
    "coala_gc_create": (_vp, [_u64]),
    "coala_gc_destroy": (None, [_vp]),
    "coala_gc_set_adj_csc": (None, [_vp, _u64, _u64]),
    "coala_gc_set_color_buffer": (None, [_vp, _u64]),
    "coala_gc_set_topk_color_buffer": (None, [_vp, _u64]),
    "coala_gc_set_topk_affinity_buffer": (None, [_vp, _u64]),
    "coala_gc_cpu_color_graph": (None, [_vp]),
    "coala_gc_cpu_color_graph_optimized": (None, [_vp, _u64, _u64]),
    "coala_gc_cpu_count_nearest_color": (None, [_vp]),
    "coala_gc_cpu_count_nearest_color_less_memory": (None, [_vp]),
    "coala_gc_cpu_calculate_color_affinity": (None, [_vp]),
    "coala_gc_get_num_color": (_u64, [_vp]),
    "coala_gc_get_num_color_node": (_u64, [_vp]),
    "coala_gc_gpu_color_graph_optimized": (_i32, [_vp, _u64, _u64, _i32]),
    "coala_gc_gpu_calculate_color_affinity": (_i32, [_vp, _i32]),
    "coala_gc_get_hops": (_i32, [_vp, _u64]),
    "coala_synth_fill_features": (_i32, [_u64, _u64, _i64, _i64, _i32, _i64, _vp]),
    "coala_synth_check_rows": (_i32, [_u64, _u64, _i64, _i32, _i64, _u64, _vp]),
    "coala_probe_read_bandwidth": (_i32, [_u64, _i64, _i32, _vp]),
    "coala_enable_peer_access": (_i32, [_i32, _i32]),
    "coala_probe_gather": (_i32, [_u64, _u64, _u64, _i64, _i32, _i32, _i32, _i32, _vp]),
    "coala_probe_gather_all": (_i32, [_i32, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _vp]),
    "coala_probe_random_read": (_i32, [_u64, _i64, _i32, _i32, _i32, _i32, _vp]),
}

_lib = None


class CoalaError(RuntimeError):
    """A failing C-ABI call (the reference surfaces std::runtime_error as RuntimeError too)."""


def lib():
    """Load (building in-tree if needed) libcoala_b200.so. Raises if that is impossible."""
    global _lib
    if _lib is None:
        path = _build.LIB
        if _build.needs_build():
            if os.path.exists(_build.NVCC):
                path = _build.build()  # re-checks under an inter-process lock
            elif not os.path.exists(path):
                raise CoalaError("libcoala_b200.so is not built and nvcc is unavailable; there is no CPU fallback")
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header/library mismatch: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def last_error():
    return lib().coala_last_error().decode(errors="replace")


def check(rc, what=""):
    if rc != 0:
        raise CoalaError("%s: %s" % (what or "libcoala_b200", last_error()))


def check_handle(h, what=""):
    if not h:
        raise CoalaError("%s: %s" % (what or "libcoala_b200", last_error()))
    return h
