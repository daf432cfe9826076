"""World-size-2/4 gloo tests (CPU) of the host-side topology and collectives that replace the
reference's MPI / nested process groups (S/Shared_Tensor.py:24-112)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, gpd, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), COALA_GPUS_PER_DOMAIN=str(gpd))
    from coala_gnn_b200.comm import MPI_Comm_Manager
    comm = MPI_Comm_Manager(-1)
    comm.initialize_nested_process_group("isolated")
    D = world // gpd
    topo = dict(node=comm.node_id, lr=comm.local_rank, ls=comm.local_size, master=comm.is_master, nm=comm.num_master_process,
                mi=comm.master_process_index, mid=comm.master_process_id, masters=comm.master_process_list,
                locals=comm.dist_local_rank_list)
    # cache colour metadata: SUM inside the domain, all-gather across domain masters
    meta = torch.full((6,), rank + 1, dtype=torch.int32)
    gathered = [torch.zeros(6, dtype=torch.int32) for _ in range(D)]
    comm.gather_cache_meta(meta, gathered)
    # seed broadcast from the domain master
    seeds = torch.arange(8, dtype=torch.int64) + 100 * rank if comm.is_master else torch.zeros(8, dtype=torch.int64)
    comm.broadcast_training_nodes(seeds)
    q.put((rank, topo, [g.tolist() for g in gathered], seeds.tolist()))
    comm.destroy_process_group()


@pytest.mark.parametrize("world,gpd", [(2, 2), (2, 1), (4, 2)])
def test_domain_topology_and_collectives(world, gpd):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, gpd, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    D = world // gpd
    for rank, topo, gathered, seeds in res:
        node = rank // gpd
        assert topo["node"] == node and topo["lr"] == rank % gpd and topo["ls"] == gpd
        assert topo["master"] == (rank % gpd == 0) and topo["nm"] == D and topo["mi"] == node
        assert topo["masters"] == [d * gpd for d in range(D)] and topo["mid"] == node * gpd
        assert topo["locals"] == list(range(node * gpd, (node + 1) * gpd))
        assert seeds == [100 * node * gpd + i for i in range(8)]
        for d in range(D):  # every rank of every domain ends up with all domains' histograms
            assert gathered[d] == [sum(r + 1 for r in range(d * gpd, (d + 1) * gpd))] * 6
