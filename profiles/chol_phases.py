#!/usr/bin/env python
"""Phase profile of the cluster Cholesky (chol_cluster_kernel): reads the clock64 stamps
CTA 0 leaves behind the tile scratch (gn_solve.cu, `stamps`) after one GN iteration and
prints the time spent per phase (panel / trailing update incl. look-ahead diagonal per
tile column, then the back substitution).
Usage: python profiles/chol_phases.py [--kf loop64] [--hw 384x512]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--kf", default="loop64")
ap.add_argument("--hw", default="96x128")
args = ap.parse_args()
H, W = (int(x) for x in args.hw.split("x"))

from frodo_slam_b200 import backends, dist as fdist, synth  # noqa: E402

dev = torch.device("cuda:0")
gp = synth.GN_PARAMS
K, und = synth.graph_edges(args.kf)
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
Twc = g["Twc"].clone()
sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                              g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                              g["Q"].view(E, N), 0, E, 1)
n = 7 * (K - 1)
T, Tc = (n + 1 + 31) // 32, (n + 31) // 32
off = (T * (T + 1) // 2 * 1024 + 32) * 8
sg.ws[off:off + 128 * 8].zero_()
for _ in range(3):
    Twc.copy_(g["Twc"])
    sg.solve(1, 0.0)
torch.cuda.synchronize()
st = sg.ws[off:off + 128 * 8].view(torch.int64).cpu().numpy().astype("float64")
mhz = 1965.0
us = lambda a, b: (b - a) / mhz  # noqa: E731
print(f"n={n} T={T} Tc={Tc}  (clock64 cycles at {mhz:.0f} MHz -> us)")
print(f"load tiles + first diagonal {us(st[0], st[1]):8.2f} us")
tot = {"panel": 0.0, "update": 0.0, "barrier_x": 0.0}
for k in range(Tc):
    a, b, c = st[2 + 3 * k], st[3 + 3 * k], st[4 + 3 * k]
    nxt = st[2 + 3 * (k + 1)] if k + 1 < Tc else st[2 + 3 * Tc]
    tot["panel"] += us(a, b)
    tot["update"] += us(b, c)
    tot["barrier_x"] += us(c, nxt)
    print(f"col {k:2d}: panel(+Y) {us(a, b):7.2f}  update(+look-ahead diag) {us(b, c):7.2f}  "
          f"barrier X {us(c, nxt):6.2f}")
print({k: round(float(v), 2) for k, v in tot.items()})
print(f"back substitution   {us(st[2 + 3 * Tc], st[3 + 3 * Tc]):8.2f} us")
print(f"total               {us(st[0], st[3 + 3 * Tc]):8.2f} us")
if st[116] > 0:
    print(f"pose update (dx, retraction)  {us(st[3 + 3 * Tc], st[116]):8.2f} us")
    wall = (st[119] - st[118]) / 1e3
    print(f"wall clock (globaltimer) start -> end of kernel body {wall:8.2f} us  => SM clock "
          f"{(st[116] - st[0]) / wall:7.1f} MHz during the kernel")
# cycles accumulated by every diagonal-tile call since the workspace was allocated (3 solves)
if st[120] > 0:
    calls = 3 * Tc
    print(f"diagonal tile: factor {st[120] / mhz / calls:6.2f} us, inverse W {st[121] / mhz / calls:6.2f} us "
          f"per tile (mean of {calls} calls)")
