"""Host-side logic of the edge-sharded Gauss-Newton (frodo-slam_b200/dist.py) on CPU:
world_size-2 `gloo` run in which every rank assembles the dense fp64 [H | g] of ITS edge
slice (per-edge blocks from the CPU oracle -- test infrastructure), the systems are summed
with dist.all_reduce exactly as ShardedGaussNewton.solve does, and the result must equal
the single-rank system; both ranks must then hold bit-identical systems (which is what lets
every rank run the same Cholesky without a broadcast of dx)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import GP, graph_inputs


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_system(g, lo, hi, ie, je, num_fix):
    from oracle import oracle
    Hs, gs = oracle.edge_blocks("rays", g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(),
                                ie[lo:hi], je[lo:hi], g["idx_ii2jj"].numpy()[lo:hi],
                                g["valid_match"].numpy()[lo:hi], g["Q"].numpy()[lo:hi],
                                GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"])
    H, gv = oracle.assemble(Hs, gs, ie[lo:hi], je[lo:hi], g["Xs"].shape[0], num_fix,
                            round_f32=False)
    return np.concatenate([H.reshape(-1), gv.reshape(-1)])


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from frodo_slam_b200 import dist as fdist
        from oracle import oracle
        g = graph_inputs("window8", 24, 32)
        E = g["ii"].shape[0]
        ie, je, _ = oracle.edge_index(g["ii"].numpy(), g["jj"].numpy())  # ranks over ALL edges
        lo, hi = fdist.edge_slice(E, rank, world)
        Hg = torch.from_numpy(_rank_system(g, lo, hi, ie, je, 1))
        dist.all_reduce(Hg, op=dist.ReduceOp.SUM)
        full = torch.from_numpy(_rank_system(g, 0, E, ie, je, 1))
        rel = float((Hg - full).norm() / full.norm())
        gathered = [torch.empty_like(Hg) for _ in range(world)]
        dist.all_gather(gathered, Hg)
        same = all(torch.equal(gathered[0], t) for t in gathered)
        out.put((rank, lo, hi, rel, same))
    finally:
        dist.destroy_process_group()


def test_edge_slice_partition():
    from frodo_slam_b200 import dist as fdist
    for E in (0, 1, 7, 24, 256, 257):
        for world in (1, 2, 3, 4, 8):
            cuts = [fdist.edge_slice(E, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == E
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(180)
def test_allreduce_of_edge_shards_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=150) for _ in range(world))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert [r[1:3] for r in res] == [(0, 12), (12, 24)]
    for _, _, _, rel, same in res:
        assert rel < 1e-12  # fp64 sums of disjoint edge sets in a different order
        assert same         # identical inputs to the replicated solve on every rank


def test_peer_exchange_is_cuda_only_and_can_be_disabled(monkeypatch):
    """The peer-memory exchange (csrc/peer_reduce.cu) is only set up for CUDA devices; on other
    devices, or with FRODO_PEER_REDUCE=0, the driver keeps the collective (NCCL / gloo)."""
    import torch
    from frodo_slam_b200 import dist as fdist
    assert fdist._peer_state(100, torch.device("cpu"), None) is None
    monkeypatch.setenv("FRODO_PEER_REDUCE", "0")
    assert fdist._peer_state(100, torch.device("cuda", 0), None) is None
