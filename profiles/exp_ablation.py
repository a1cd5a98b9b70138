#!/usr/bin/env python
"""Times frodo_gn_accumulate (64-KF rays graph) with the library given in FRODO_ABL_LIB
(ablation builds of gn_accumulate.cu, -DFRODO_ABL=1 no arithmetic, 2 gather at own index,
3 both).  Development aid."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from frodo_slam_b200 import _lib  # noqa: E402

if os.environ.get("FRODO_ABL_LIB"):
    _lib.LIB_PATH = os.path.join(ROOT, "frodo-slam_b200", os.environ["FRODO_ABL_LIB"])
from frodo_slam_b200 import backends, dist as fdist, synth  # noqa: E402
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
Twc = g["Twc"].clone()
sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                              g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                              g["Q"].view(E, N), 0, E, 1)
sg.solve(1, 0.0)


def acc():
    _lib.check(lib.frodo_gn_accumulate(
        ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc, _ptr(sg.ii_row),
        _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N, E, 1, None, None, None,
        _ptr(sg.ws), sg.nbytes, _stream()))


for _ in range(3):
    acc()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    acc()
e1.record()
torch.cuda.synchronize()
print(f'{os.environ.get("FRODO_ABL_LIB", "libfrodo_b200.so"):24s} '
      f'{e0.elapsed_time(e1) / 20 * 1e3:8.1f} us per launch')
