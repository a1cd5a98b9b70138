#!/usr/bin/env python
"""Runs every kernel of the hot path a few times at bench sizes, for ncu captures:
  GN rays on the 64-KF graph (2 iterations), iter_proj + refine_matches on 8 pairs at 512x384.
Usage: python profiles/prof_ops.py [--kf loop64] [--pairs 8]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--kf", default="loop64")
ap.add_argument("--pairs", type=int, default=8)
ap.add_argument("--iters", type=int, default=2)
args = ap.parse_args()

import mast3r_slam_backends as be  # noqa: E402
from frodo_slam_b200 import synth  # noqa: E402

dev = torch.device("cuda:0")
H, W = 384, 512
gp, mp = synth.GN_PARAMS, synth.MATCH_PARAMS
K, und = synth.graph_edges(args.kf)
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
for _ in range(2):
    T = g["Twc"].clone()
    be.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"],
                         g["Q"], gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], 1,
                         args.iters, 0.0)
torch.cuda.synchronize()
del g
B = args.pairs
d = synth.matching_pairs(B, H, W, seed=0, device=dev)
rays, pts, p0 = synth.prep_for_iter_proj(d["X11"], d["X21"])
D11 = d["D11"].half().contiguous()
D21 = d["D21"].half().reshape(B, H * W, -1).contiguous()
for _ in range(2):
    p, conv = be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"],
                           mp["convergence_thresh"])
    be.refine_matches(D11, D21, p.long(), mp["radius"], mp["dilation_max"])
torch.cuda.synchronize()
# the matching pipeline in one call (prep, occlusion, casts, finish) and the quality glue
from frodo_slam_b200 import matching as fm, tracking  # noqa: E402
X11, X21 = d["X11"].float().contiguous(), d["X21"].float().contiguous()
D11f, D21f = d["D11"].float().contiguous(), d["D21"].float().contiguous()
idx, valid = fm.match(X11, X21, D11f, D21f, cfg=mp)
Qs = torch.rand(B, H * W, 1, device=dev) * 3
fm.match_quality(Qs, Qs.clone(), idx, valid.reshape(B, H * W, 1), 1.5, Qs, Qs, 0.0)
torch.cuda.synchronize()
del d, X11, X21, D11f, D21f, D11, D21
# the tracker's pose solve
c = synth.tracking_case(H, W, seed=7)
t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in c.items()}
for _ in range(2):
    tracking.track_pose("rays", t["Xf_all"], t["Xk"], t["T_WCf"], t["T_WCk"], t["Qk"], t["valid"],
                        idx_f2k=t["idx_f2k"], cfg=synth.TRACK_PARAMS)
torch.cuda.synchronize()
print("prof_ops done")
