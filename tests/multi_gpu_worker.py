"""Worker of tests/test_gpu_multi.py: run under torchrun with >= 2 GPUs.  Every rank solves the
same keyframe graphs (8-keyframe window and the 64-keyframe loop-closure graph of
BASELINE.json configs[2]) with the edge-sharded driver, with each of its three exchange paths --
per-edge blocks pushed to the peers from inside the accumulate kernel (one C-ABI call per
solve, frodo_gauss_newton_sharded), the dense peer-memory all-reduce kernel
(csrc/peer_reduce.cu), NCCL's all-reduce -- and checks
  * poses bit-identical on all ranks (every path),
  * the paths against each other: same poses to 1e-6 (the sums are ordered differently),
  * sharded vs the single-GPU operator: 1e-5,
  * odometry terms through the sharded driver == single-GPU gauss_newton_rays_odom to 1e-5."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from frodo_slam_b200 import backends, dist as fdist, synth  # noqa: E402
import mast3r_slam_backends as be  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    gp, op = synth.GN_PARAMS, synth.ODOM_PARAMS
    H, W = 48, 64
    for cfg in ("window8", "loop64"):
        run_graph(cfg, H, W, rank, world, dev, gp, op)
    dist.barrier()
    dist.destroy_process_group()


def run_graph(cfg, H, W, rank, world, dev, gp, op):
    K, und = synth.graph_edges(cfg)
    g = synth.keyframe_graph(K, H, W, und, seed=5, device="cpu")
    g = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    od = synth.odometry_edges(g, device=dev)
    E, N = g["ii"].shape[0], H * W
    lo, hi = fdist.edge_slice(E, rank, world)
    prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                               C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])

    def sharded(odom):
        T = g["Twc"].clone()
        sg = fdist.ShardedGaussNewton(prm, T, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                      g["idx_ii2jj"].view(E, N)[lo:hi].contiguous(),
                                      g["valid_match"].view(E, N)[lo:hi].contiguous(),
                                      g["Q"].view(E, N)[lo:hi].contiguous(), lo, hi, 1)
        if odom:
            sg.set_odometry(od["odom_ii"], od["odom_jj"], od["Tij"], od["s_bar"],
                            op["sigma_odom_t"], op["sigma_odom_r"], op["sigma_scale_prior"])
        for _ in range(3):  # several solves: epochs and buffer parity keep advancing
            T.copy_(g["Twc"])
            sg.solve(4, 0.0)
        torch.cuda.synchronize()
        assert int(sg.flags[1].item()) == 4 and int(sg.flags[2].item()) == 0
        return T, ("blocks" if sg.blocks is not None else "peer" if sg.peer is not None else "nccl")

    def with_env(**kv):
        for k, v in kv.items():
            os.environ[k] = v
        fdist._PEER_CACHE.clear()

    def same_on_all_ranks(T):
        out = [torch.empty_like(T) for _ in range(world)]
        dist.all_gather(out, T)
        return all(torch.equal(out[0], o) for o in out)

    with_env(FRODO_BLOCK_EXCHANGE="1", FRODO_PEER_REDUCE="1")
    T_blk, used_blk = sharded(False)
    assert same_on_all_ranks(T_blk), "block exchange: poses differ across ranks"
    with_env(FRODO_BLOCK_EXCHANGE="0")
    T_peer, used_peer = sharded(False)
    assert used_peer != "blocks"
    assert same_on_all_ranks(T_peer), "peer all-reduce: poses differ across ranks"
    with_env(FRODO_PEER_REDUCE="0")
    T_nccl, used2 = sharded(False)
    assert used2 == "nccl"
    assert same_on_all_ranks(T_nccl), "NCCL path: poses differ across ranks"
    with_env(FRODO_BLOCK_EXCHANGE="1", FRODO_PEER_REDUCE="1")
    d_bp = (T_blk - T_peer).abs().max().item()
    assert d_bp < 1e-6, d_bp
    T_one = g["Twc"].clone()
    be.gauss_newton_rays(T_one, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                         g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"],
                         gp["Q_conf"], 1, 4, 0.0)
    d_pn = (T_peer - T_nccl).abs().max().item()
    d_p1 = (T_peer - T_one).abs().max().item()
    assert d_pn < 1e-6 and d_p1 < 1e-5, (d_pn, d_p1)
    # odometry terms after the exchange
    T_od, _ = sharded(True)
    assert same_on_all_ranks(T_od)
    T_od1 = g["Twc"].clone()
    be.gauss_newton_rays_odom(T_od1, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                              g["valid_match"], g["Q"], od["odom_ii"], od["odom_jj"], od["Tij"],
                              od["s_bar"], op["sigma_odom_t"], op["sigma_odom_r"], gp["sigma_ray"],
                              gp["sigma_dist"], op["sigma_scale_prior"], gp["C_conf"], gp["Q_conf"],
                              1, 4, 0.0)
    d_od = (T_od - T_od1).abs().max().item()
    assert d_od < 1e-5, d_od
    if rank == 0:
        print(f"MULTI_GPU_OK {cfg} world={world} exchange={used_blk}/{used_peer} "
              f"blocks_vs_peer={d_bp:.2e} peer_vs_nccl={d_pn:.2e} sharded_vs_single={d_p1:.2e} "
              f"odom={d_od:.2e}", flush=True)


if __name__ == "__main__":
    main()
