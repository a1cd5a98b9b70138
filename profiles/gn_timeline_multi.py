#!/usr/bin/env python
"""Where one Gauss-Newton iteration of the edge-sharded solve goes at N GPUs (run under torchrun,
FRODO_CHOL_STAMPS=1): per rank the accumulate kernel alone, the pack, the Cholesky alone (dense
input, no wait), the whole 10-iteration solve, and the time worker 0 of the Cholesky spent
waiting for the peers' edge blocks (clock64 stamps).  What is left is launch gaps + rank skew
not visible as waiting.  Usage:
  FRODO_CHOL_STAMPS=1 python -m torch.distributed.run --nproc-per-node N profiles/gn_timeline_multi.py"""
import ctypes
import json
import os
import sys

os.environ.setdefault("FRODO_CHOL_STAMPS", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from frodo_slam_b200 import _lib, backends, dist as fdist, synth  # noqa: E402
from frodo_slam_b200.backends import _ptr, _stream  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
H, W, iters = 384, 512, 10
gp = synth.GN_PARAMS
K, und = synth.graph_edges("loop64")
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
lo, hi = fdist.edge_slice(E, rank, world)
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
idx, valid, Q = (g[k].view(E, N)[lo:hi].contiguous() for k in ("idx_ii2jj", "valid_match", "Q"))
Twc0 = g["Twc"].clone()
Twc = g["Twc"].clone()
sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"], idx, valid, Q,
                              lo, hi, 1)
lib = _lib.load()


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n  # us


def solve():
    Twc.copy_(Twc0)
    sg.solve(iters, 0.0)


us_solve = timed(solve, 10)
st_off = lib.frodo_gn_chol_stamps_offset(K)
stamps = sg.ws_all[st_off:st_off + 1024 * 8].view(torch.int64)
stamps[900:902].zero_()
us_solve = timed(solve, 10, warm=0)
wait_cyc, wait_n = (int(x) for x in stamps[900:902].tolist())


def acc_only():
    _lib.check(lib.frodo_gn_accumulate(
        ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc, _ptr(sg.ii_row),
        _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N, sg.E_local, 1, None, None,
        None, _ptr(sg.ws), sg.nbytes, _stream()))


def pack_only():
    _lib.check(lib.frodo_gn_pack_points(_ptr(g["Xs"]), _ptr(g["Cs"]), K, N, sg.xc, _stream()))


_lib.check(lib.frodo_gn_workspace_init(_ptr(sg.ws), sg.nbytes, K, max(sg.E_local, 1), N, _stream()))
pack_only()
us_acc = timed(acc_only, 30)
us_pack = timed(pack_only, 10)
# Cholesky alone on a dense system (no wait): accumulate + assemble once, then time the solve
n = 7 * (K - 1)
Hg0 = torch.empty((n + 1) * n, dtype=torch.float64, device=dev)
ii_row = torch.empty(E, dtype=torch.int32, device=dev)
jj_row = torch.empty(E, dtype=torch.int32, device=dev)
_lib.check(lib.frodo_gn_edge_index(_ptr(g["ii"]), _ptr(g["jj"]), E, _ptr(ii_row), _ptr(jj_row), None, _stream()))
nb = lib.frodo_gn_workspace_bytes(K, E, N)
ws = torch.empty(nb, dtype=torch.uint8, device=dev)
_lib.check(lib.frodo_gn_workspace_init(_ptr(ws), nb, K, E, N, _stream()))
Twc.copy_(Twc0)
_lib.check(lib.frodo_gn_accumulate(
    ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, None, _ptr(ii_row),
    _ptr(jj_row), _ptr(g["idx_ii2jj"]), _ptr(g["valid_match"]), _ptr(g["Q"]), K, N, E, 1, _ptr(Hg0),
    None, None, _ptr(ws), nb, _stream()))
Hg = Hg0.clone()
dx = torch.zeros(K - 1, 7, device=dev)
T2 = Twc0.clone()


def chol_only():
    Hg.copy_(Hg0)
    _lib.check(lib.frodo_gn_solve_retract(_ptr(Hg), _ptr(T2), _ptr(dx), K, 1, 0.0, None, None,
                                          _ptr(ws), nb, _stream()))


def copy_only():
    Hg.copy_(Hg0)


us_chol = timed(chol_only, 20) - timed(copy_only, 20)
row = dict(rank=rank, world=world, edges=hi - lo, us_per_iteration=us_solve / iters,
           us_accumulate=us_acc, us_pack_per_iteration=us_pack / iters, us_cholesky_alone=us_chol,
           us_wait_for_peers_per_iteration=(wait_cyc / 1965.0 / max(wait_n, 1)) if wait_n else 0.0)
row["us_rest"] = row["us_per_iteration"] - us_acc - row["us_pack_per_iteration"] - us_chol \
    - row["us_wait_for_peers_per_iteration"]
if world > 1:
    rows = [None] * world
    dist.all_gather_object(rows, row)
else:
    rows = [row]
if rank == 0:
    for r in rows:
        print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in r.items()}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
