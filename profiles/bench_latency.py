#!/usr/bin/env python
"""Single-call latencies of the per-frame path at 512x384 (what one tracked frame costs):
match() for B = 1 and B = 2 pairs, its two kernels alone, match_quality, the tracker solve."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mast3r_slam_backends as be  # noqa: E402
from frodo_slam_b200 import matching as fm, synth  # noqa: E402

dev = torch.device("cuda:0")
H, W = 384, 512
mp = synth.MATCH_PARAMS


def t(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for B in [int(x) for x in os.environ.get("FRODO_LAT_B", "1,2,8").split(",")]:
    d = synth.matching_pairs(B, H, W, seed=0, device=dev)
    X11, X21 = d["X11"].float().contiguous(), d["X21"].float().contiguous()
    D11, D21 = d["D11"].float().contiguous(), d["D21"].float().contiguous()
    rays, pts, p0 = synth.prep_for_iter_proj(X11, X21)
    D11h, D21h = D11.half().contiguous(), D21.half().reshape(B, H * W, -1).contiguous()
    us_m = t(lambda: fm.match(X11, X21, D11, D21, cfg=mp))
    us_i = t(lambda: be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"], mp["convergence_thresh"]))
    p1 = be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"], mp["convergence_thresh"])[0].long()
    us_r = t(lambda: be.refine_matches(D11h, D21h, p1, mp["radius"], mp["dilation_max"]))
    print(f"B={B}: match() {us_m:7.1f} us   iter_proj {us_i:6.1f} us   refine_matches {us_r:6.1f} us "
          f"({B * H * W / us_m / 1e3:.2f} G px-matches/s through match())")
