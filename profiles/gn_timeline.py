#!/usr/bin/env python
"""Where one Gauss-Newton iteration spends its time on the stream: CUDA events between the
launches of the split C ABI (accumulate + assemble | Cholesky + retraction), 64-KF 512x384 graph."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from frodo_slam_b200 import _lib, backends, dist as fdist, synth  # noqa: E402
from frodo_slam_b200.backends import _ptr, _stream  # noqa: E402

dev = torch.device("cuda:0")
gp = synth.GN_PARAMS
K, und = synth.graph_edges(os.environ.get("FRODO_TL_GRAPH", "loop64"))
H, W = 384, 512
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
Twc = g["Twc"].clone()
sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                              g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                              g["Q"].view(E, N), 0, E, 1)
sg.solve(3, 0.0)
lib = _lib.load()
iters = 20
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * iters + 1)]
st = _stream()
Twc.copy_(g["Twc"])
sg.flags.zero_()
torch.cuda.synchronize()
ev[0].record()
for i in range(iters):
    _lib.check(lib.frodo_gn_accumulate(ctypes.byref(prm), _ptr(Twc), _ptr(sg.Xs), _ptr(sg.Cs), None,
                                       sg.xc, _ptr(sg.ii_row), _ptr(sg.jj_row), _ptr(sg.idx),
                                       _ptr(sg.valid), _ptr(sg.Q), K, N, E, 1, _ptr(sg.Hg), None,
                                       None, _ptr(sg.ws), sg.nbytes, st))
    ev[2 * i + 1].record()
    _lib.check(lib.frodo_gn_solve_retract(_ptr(sg.Hg), _ptr(Twc), _ptr(sg.dx), K, 1, 0.0, None,
                                          None, _ptr(sg.ws), sg.nbytes, st))
    ev[2 * i + 2].record()
torch.cuda.synchronize()
acc = [ev[2 * i].elapsed_time(ev[2 * i + 1]) * 1e3 for i in range(2, iters)]
sol = [ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) * 1e3 for i in range(2, iters)]
tot = ev[4].elapsed_time(ev[2 * iters]) * 1e3 / (iters - 2)
print(f"accumulate + assemble segment {sum(acc) / len(acc):7.1f} us   Cholesky + retraction segment "
      f"{sum(sol) / len(sol):7.1f} us   per iteration {tot:7.1f} us")
