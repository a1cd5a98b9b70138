#!/usr/bin/env python
"""Kernel start/end times of one gauss_newton_rays call (CUPTI through torch.profiler): how much
of an iteration is the gap between the accumulate and the Cholesky kernels."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from frodo_slam_b200 import _lib
if os.environ.get("ALTLIB"):
    _lib.LIB_PATH = os.path.join(ROOT, os.environ["ALTLIB"])
from frodo_slam_b200 import backends, synth
dev = torch.device("cuda:0")
H, W = (int(x) for x in os.environ.get("HW", "384x512").split("x"))
gp = synth.GN_PARAMS
for cfg in os.environ.get("CFGS", "window8,loop64").split(","):
    K, und = synth.graph_edges(cfg)
    g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
    def solve():
        T = g["Twc"].clone()
        backends.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                   g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"],
                                   gp["C_conf"], gp["Q_conf"], 1, 10, 0.0)
    for _ in range(3): solve()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        solve(); torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: e.time_range.start)
    t0 = ev[0].time_range.start
    print(f"== {cfg}")
    prev_end = None
    rows = []
    for e in ev:
        s, en = e.time_range.start - t0, e.time_range.end - t0
        gap = (s - prev_end) if prev_end is not None else 0.0
        rows.append((e.name[:48], s, en - s, gap))
        prev_end = en
    for r in rows[:14]: print(f"  {r[0]:50s} start {r[1]:9.1f} dur {r[2]:8.1f} gap {r[3]:6.1f}")
    import collections
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for n, s, d, gp_ in rows:
        a = agg[n]; a[0] += 1; a[1] += d; a[2] += gp_
    for n, (c, d, gp_) in agg.items():
        print(f"  {n:50s} n={c:3d} mean dur {d/c:8.1f} mean gap before {gp_/c:6.1f}")
    print(f"  total span {rows[-1][1] + rows[-1][2]:.1f} us")
    del g
