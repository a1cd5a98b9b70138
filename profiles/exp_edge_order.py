#!/usr/bin/env python
"""Experiment: does the order in which edges are processed change the accumulate time
(L2 reuse of keyframe pointmaps between co-resident CTAs)?  Times frodo_gn_accumulate on the
64-KF graph with the edges in caller order and sorted by (min(i,j), max(i,j))."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from frodo_slam_b200 import _lib, backends, dist as fdist, synth  # noqa: E402
from frodo_slam_b200.backends import _ptr, _stream  # noqa: E402

dev = torch.device("cuda:0")
H, W = 384, 512
N = H * W
gp = synth.GN_PARAMS
K, und = synth.graph_edges("loop64")
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E = g["ii"].shape[0]
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
lib = _lib.load()


def run(order, label):
    o = torch.as_tensor(order, device=dev)
    ii, jj = g["ii"][o].contiguous(), g["jj"][o].contiguous()
    idx = g["idx_ii2jj"].view(E, N)[o].contiguous()
    valid = g["valid_match"].view(E, N)[o].contiguous()
    Q = g["Q"].view(E, N)[o].contiguous()
    Twc = g["Twc"].clone()
    sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, ii, jj, idx, valid, Q, 0, E, 1)
    sg.solve(1, 0.0)

    def acc():
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc,
            _ptr(sg.ii_row), _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N, E, 1,
            None, None, None, _ptr(sg.ws), sg.nbytes, _stream()))
    for _ in range(3):
        acc()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        acc()
    e1.record()
    torch.cuda.synchronize()
    print(f"{label:28s} {e0.elapsed_time(e1) / 20 * 1e3:8.1f} us per launch")
    del sg


ii, jj = g["ii"].tolist(), g["jj"].tolist()
run(list(range(E)), "caller order")
run(sorted(range(E), key=lambda e: (min(ii[e], jj[e]), max(ii[e], jj[e]), ii[e])), "sorted (min,max)")
run(sorted(range(E), key=lambda e: (ii[e], jj[e])), "sorted by i")
run(sorted(range(E), key=lambda e: (jj[e], ii[e])), "sorted by j")
