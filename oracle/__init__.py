"""TEST INFRASTRUCTURE ONLY — CPU oracle for the COALA-GNN mini-batch data path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product (``coala_gnn_b200``)
never imports it and has no CPU fallback.

Two libraries sit behind it:

* ``oracle/_build/liboracle.so`` — this repo's plain-C restatement (``oracle/c/*.c``);
* ``oracle/_ref/libcoala_ref.so`` — the reference's own ``node_distributor_pybind.cuh`` and
  ``graph_coloring.cpp`` compiled from ``/root/reference`` by ``oracle/Makefile`` (prebuilt
  file travels to the GPU box; the sources do not).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "libcoala_ref.so")

_u64, _i64, _i32, _vp = C.c_uint64, C.c_int64, C.c_int, C.c_void_p


def build(quiet=True):
    """Compile the C restatement and, if /root/reference is present, oracle/_ref."""
    out = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.orc_mix64.restype = _u64
        L.orc_mix64.argtypes = [_u64]
        L.orc_feat_value.restype = C.c_float
        L.orc_feat_value.argtypes = [_u64, _u64, _u64]
        L.orc_fill_features.argtypes = [_vp, _u64, _u64, _u64, _u64, _u64]
        L.orc_distribute.argtypes = [_vp, _i64, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp]
        L.orc_cache_create.restype = _vp
        L.orc_cache_create.argtypes = [_u64, C.c_uint32, _i32, _i32, _vp, _i32, _i32]
        L.orc_cache_destroy.argtypes = [_vp]
        L.orc_cache_num_sets.restype = _u64
        L.orc_cache_num_sets.argtypes = [_vp]
        L.orc_cache_read.argtypes = [_vp, _vp, _i64, _vp, _vp, _i64]
        L.orc_cache_stats.argtypes = [_vp, _vp, _vp]
        for f in ("orc_cache_keys", "orc_cache_color_meta", "orc_cache_color_counters", "orc_cache_set_cnt"):
            getattr(L, f).restype = _vp
            getattr(L, f).argtypes = [_vp, _i32]
        L.orc_sample_layer.restype = _i64
        L.orc_sample_layer.argtypes = [_vp, _vp, _vp, _vp, _i64, _i32, _u64, _i64, _vp, _vp, _vp]
        L.orc_to_block.restype = _i64
        L.orc_to_block.argtypes = [_vp, _i64, _vp, _i64, _vp, _vp]
        L.orc_sample_blocks.restype = _i64
        L.orc_sample_blocks.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _i32, _u64, _i64, _vp, _vp, _i32]
        L.orc_index_select.argtypes = [_vp, _i64, _vp, _i64, _vp, _i32]
        _lib = L
    return _lib


def have_ref():
    return os.path.exists(_REF)


def ref():
    """The reference's own compiled host code (kind = "reference")."""
    global _ref
    if _ref is None:
        if not os.path.exists(_REF):
            raise RuntimeError("oracle/_ref/libcoala_ref.so missing: run `make -C oracle` where /root/reference exists")
        R = C.CDLL(_REF)
        R.ref_srand.argtypes = [C.c_uint]
        R.ref_nd_create.restype = _vp
        R.ref_nd_create.argtypes = [_u64, _i32, _i32, _i32, _i32, C.c_char_p, C.c_char_p, C.c_char_p]
        R.ref_nd_distribute.argtypes = [_vp, _u64, _u64, _vp, _i32]
        R.ref_nd_num_colors.argtypes = [_vp]
        R.ref_nd_destroy.argtypes = [_vp]
        R.ref_gc_create.restype = _vp
        R.ref_gc_create.argtypes = [_u64]
        for f in ("ref_gc_destroy", "ref_gc_color_graph", "ref_gc_count_nearest_color",
                  "ref_gc_count_nearest_color_less_memory", "ref_gc_calculate_color_affinity"):
            getattr(R, f).argtypes = [_vp]
        R.ref_gc_set_adj_csc.argtypes = [_vp, _u64, _u64]
        for f in ("ref_gc_set_color_buffer", "ref_gc_set_topk_color_buffer", "ref_gc_set_topk_affinity_buffer"):
            getattr(R, f).argtypes = [_vp, _u64]
        R.ref_gc_color_graph_optimized.argtypes = [_vp, _u64, _u64]
        R.ref_gc_get_num_color.restype = _u64
        R.ref_gc_get_num_color.argtypes = [_vp]
        R.ref_gc_get_num_color_node.restype = _u64
        R.ref_gc_get_num_color_node.argtypes = [_vp]
        _ref = R
    return _ref


def _p(a):
    return a.ctypes.data_as(_vp) if a is not None else None


from .api import *  # noqa: E402,F401
