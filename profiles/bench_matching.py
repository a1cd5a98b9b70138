#!/usr/bin/env python
"""CUDA-event timing of iter_proj and refine_matches alone (B pairs at 512x384, inputs in HBM,
> L2), plus displacement statistics of the synthetic matches.  Development aid; bench.py's
"matching" object is the reported number.
Usage: python profiles/bench_matching.py [--pairs 32] [--reps 10] [--ops both|iter|refine]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=32)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--ops", default="both")
args = ap.parse_args()

import mast3r_slam_backends as be  # noqa: E402
from frodo_slam_b200 import synth  # noqa: E402

dev = torch.device("cuda:0")
H, W = 384, 512
mp = synth.MATCH_PARAMS
B = args.pairs
base = min(B, 8)
d = synth.matching_pairs(base, H, W, seed=0, device=dev)
rep = (B + base - 1) // base
tile = lambda x: x.repeat((rep,) + (1,) * (x.dim() - 1))[:B].contiguous()  # noqa: E731
rays, pts, p0 = synth.prep_for_iter_proj(tile(d["X11"]), tile(d["X21"]))
D11 = tile(d["D11"]).half().contiguous()
D21 = tile(d["D21"]).half().reshape(B, H * W, -1).contiguous()


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / args.reps


p, conv = be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"], mp["convergence_thresh"])
p1 = p.long()
px = B * H * W
if args.ops in ("both", "iter"):
    ms = timed(lambda: be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"],
                                    mp["convergence_thresh"]))
    disp = (p - p0).norm(dim=-1)
    print(f"iter_proj      {ms:8.3f} ms  {px / ms / 1e6:8.2f} Gpx/s  ({ms / B * 1e3:.1f} us/pair)  "
          f"converged {conv.float().mean().item():.3f}  |p-p0| mean {disp.mean().item():.1f} "
          f"max {disp.max().item():.1f} px")
if args.ops in ("both", "refine"):
    ms = timed(lambda: be.refine_matches(D11, D21, p1, mp["radius"], mp["dilation_max"]))
    pn, sc = be.refine_matches(D11, D21, p1, mp["radius"], mp["dilation_max"])
    mv = (pn - p1).abs().amax(dim=-1).float()
    print(f"refine_matches {ms:8.3f} ms  {px / ms / 1e6:8.2f} Gpx/s  ({ms / B * 1e3:.1f} us/pair)  "
          f"|move| mean {mv.mean().item():.2f} max {mv.max().item():.0f} px")
