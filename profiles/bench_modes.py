#!/usr/bin/env python
"""CUDA-event timing of the three Gauss-Newton modes on the 64-keyframe bench graph
(10 iterations, delta_thresh = 0) and of their accumulate launch alone.  Development aid."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import mast3r_slam_backends as be  # noqa: E402
from frodo_slam_b200 import _lib, backends, dist as fdist, synth  # noqa: E402
from frodo_slam_b200.backends import _ptr, _stream  # noqa: E402

dev = torch.device("cuda:0")
H, W = 384, 512
N = H * W
gp = synth.GN_PARAMS
K, und = synth.graph_edges("loop64")
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E = g["ii"].shape[0]
Kmat = synth.tum_intrinsics(H, W).to(dev)
lib = _lib.load()


def timed(fn, n=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


a = (g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"])
calls = {
    "rays": lambda T: be.gauss_newton_rays(T, *a, gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"],
                                           gp["Q_conf"], 1, 10, 0.0),
    "points": lambda T: be.gauss_newton_points(T, *a, gp["sigma_point"], gp["C_conf"], gp["Q_conf"],
                                               10, 0.0),
    "calib": lambda T: be.gauss_newton_calib(T, g["Xs"], g["Cs"], Kmat, *a[2:], H, W,
                                             gp["pixel_border"], gp["depth_eps"], gp["sigma_pixel"],
                                             gp["sigma_depth"], gp["C_conf"], gp["Q_conf"], 10, 0.0),
}
for mode, fn in calls.items():
    ms = timed(lambda: fn(g["Twc"].clone()))
    if mode == "rays":
        prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                                   C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
    elif mode == "points":
        prm = backends.make_params("points", sigma_point=gp["sigma_point"], C_thresh=gp["C_conf"],
                                   Q_thresh=gp["Q_conf"])
    else:
        prm = backends.make_params("calib", height=H, width=W, pixel_border=gp["pixel_border"],
                                   z_eps=gp["depth_eps"], sigma_pixel=gp["sigma_pixel"],
                                   sigma_depth=gp["sigma_depth"], C_thresh=gp["C_conf"],
                                   Q_thresh=gp["Q_conf"])
    Twc = g["Twc"].clone()
    sg = fdist.ShardedGaussNewton(prm, Twc, g["Xs"], g["Cs"], Kmat if mode == "calib" else None,
                                  g["ii"], g["jj"], g["idx_ii2jj"].view(E, N),
                                  g["valid_match"].view(E, N), g["Q"].view(E, N), 0, E, 1)
    sg.solve(1, 0.0)

    def acc():
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]),
            _ptr(Kmat) if mode == "calib" else None, sg.xc, _ptr(sg.ii_row), _ptr(sg.jj_row),
            _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N, E, 1, None, None, None, _ptr(sg.ws),
            sg.nbytes, _stream()))

    us = timed(acc, 10) * 1e3
    print(f"{mode:7s} 10 iterations {ms:7.2f} ms = {1e4 / ms:7.1f} iters/s   accumulate {us:7.1f} us "
          f"= {45 * N * E / us / 1e3:6.0f} GB/s algorithmic")
