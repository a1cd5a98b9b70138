#!/usr/bin/env python
"""How much of chol_cluster_kernel's duration is launch / teardown: the same launch with the
`done` flag preset (the kernel returns at its first instruction) next to a real solve.
Run under `ncu --metrics gpu__time_duration.sum -k regex:chol`."""
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
K, und = synth.graph_edges("loop64")
H, W = 96, 128
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
Twc = g["Twc"].clone()
sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                              g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                              g["Q"].view(E, N), 0, E, 1)
sg.solve(2, 0.0)  # two real solves
lib = _lib.load()
sg.flags[0] = 1   # done: the next launches return immediately
done = ctypes.c_void_p(sg.flags.data_ptr())
for _ in range(3):
    _lib.check(lib.frodo_gn_solve_retract(_ptr(sg.Hg), _ptr(Twc), _ptr(sg.dx), K, 1, 0.0, done,
                                          None, _ptr(sg.ws), sg.nbytes, _stream()))
torch.cuda.synchronize()
print("done")
