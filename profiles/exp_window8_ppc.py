#!/usr/bin/env python
"""Accumulate kernel alone on the 8-keyframe window (E = 24) and on the 32-keyframe graph
(E = 128) for the point span per CTA given in FRODO_GN_PPC (experiment)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from frodo_slam_b200 import _lib, backends, dist as fdist, synth
from frodo_slam_b200.backends import _ptr, _stream
dev = torch.device("cuda:0")
H, W = 384, 512
gp = synth.GN_PARAMS
lib = _lib.load()
for cfg in os.environ.get("CFGS", "window8,calib32").split(","):
    K, und = synth.graph_edges(cfg)
    g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
    E, N = g["ii"].shape[0], H * W
    prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                               C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
    sg = fdist.ShardedGaussNewton(prm, g["Twc"].clone(), g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                  g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                                  g["Q"].view(E, N), 0, E, 1)
    _lib.check(lib.frodo_gn_workspace_init(_ptr(sg.ws), sg.nbytes, K, E, N, _stream()))
    _lib.check(lib.frodo_gn_pack_points(_ptr(g["Xs"]), _ptr(g["Cs"]), K, N, sg.xc, _stream()))
    def acc():
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(prm), _ptr(sg.Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc,
            _ptr(sg.ii_row), _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N, E, 1,
            None, None, None, _ptr(sg.ws), sg.nbytes, _stream()))
    for _ in range(3): acc()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30): acc()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 30
    print(f"{cfg}: E={E} ppc={os.environ.get('FRODO_GN_PPC','auto')} accumulate {us:7.1f} us = "
          f"{45 * N * E / us / 1e3:6.0f} GB/s")
    del sg, g
