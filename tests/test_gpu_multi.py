"""N > 1 on real GPUs: launches tests/multi_gpu_check.py under torchrun (one process per GPU) when the box has
at least two GPUs; skipped on a single-GPU box (the world-size-2 host logic is covered on CPU by
tests/test_comm_gloo.py).  The script checks every batch of every rank bit-exactly (see its docstring)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(nproc, extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(HERE, "multi_gpu_check.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "multi_gpu_check: all good" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def test_one_domain_two_gpus_all_backends():
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    out = _run(2, [])
    for backend in ("nvshmem", "isolated", "nccl"):
        assert ("ok %-8s" % backend) in out


def test_two_domains_of_one_gpu_isolated():
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--gpus_per_domain", "1", "--backends", "isolated,nvshmem"])


def _run_stress(nproc, extra):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(HERE, "multi_gpu_stress.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "multi_gpu_stress: all good" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_eviction_storm_one_gpu():
    """The storm of tests/multi_gpu_stress.py on one GPU (device-scope protocol; runs on the driver's 1-GPU box too)."""
    _run_stress(1, ["--launches", "2000"])


def test_eviction_storm_across_gpus():
    """Cross-GPU seqlock / remote-fill stress: every GPU of the box hammers the same sets of every shard (system scope)."""
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    _run_stress(min(n, 8), ["--launches", "10000"])
