"""Which binding of the C ABI the COALA_GNN package drives.

Two bindings of include/coala_b200.h exist, with the same seven classes as the reference's module
(M/COALA_GNN_Pybind.cu:27-79):
  * coala_gnn_b200/pybind.py — ctypes (default; drops the GIL around every call);
  * COALA_GNN_Pybind/_native*.so — the compiled pybind11 module built from csrc_pybind/COALA_GNN_Pybind.cpp, i.e.
    what a maintainer of the reference would compile in place of COALA_GNN_Pybind.cu.
`COALA_PYBIND_NATIVE=1` (or use_native(True) before the objects are built) selects the compiled one."""
import os

from . import pybind as _ct

_native = None


def use_native(flag=True):
    global _native
    if flag:
        from . import build as _build
        _build.build_pybind()
        import COALA_GNN_Pybind._native as m
        _native = m
    else:
        _native = None


def is_native():
    return _native is not None


def mod():
    return _native if _native is not None else _ct


def make_shared_uva(path, size, node_id, local_comm, device, zero):
    """SharedUVAManager with the create/open order of M/shared_UVA.cuh:65-81 (local rank 0 creates, the others open
    after the barrier)."""
    if _native is None:
        return _ct.SharedUVAManager(path, size, node_id, 0, local_comm, device=device, zero=zero)
    rank = local_comm.rank() if local_comm is not None else 0
    if rank == 0:
        h = _native.SharedUVAManager(path, int(size), int(node_id), 0, 1, int(device), bool(zero))
        if local_comm is not None:
            local_comm.barrier()
        return h
    local_comm.barrier()
    return _native.SharedUVAManager(path, int(size), int(node_id), 0, 0, int(device), False)


if os.environ.get("COALA_PYBIND_NATIVE", "0") == "1":
    use_native(True)
