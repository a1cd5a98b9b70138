#!/usr/bin/env python
"""Accumulate-kernel time of each 32-edge slice of the 64-keyframe graph on ONE GPU: tells data
effects (gather locality of a slice) from per-GPU effects in the 8-GPU timeline."""
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
gp = synth.GN_PARAMS
K, und = synth.graph_edges("loop64")
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
lib = _lib.load()
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for r in range(world):
    lo, hi = fdist.edge_slice(E, r, world)
    idx, valid, Q = (g[k].view(E, N)[lo:hi].contiguous() for k in ("idx_ii2jj", "valid_match", "Q"))
    sg = fdist.ShardedGaussNewton(prm, g["Twc"].clone(), g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                  idx, valid, Q, lo, hi, 1)
    _lib.check(lib.frodo_gn_workspace_init(_ptr(sg.ws), sg.nbytes, K, max(sg.E_local, 1), N, _stream()))
    _lib.check(lib.frodo_gn_pack_points(_ptr(g["Xs"]), _ptr(g["Cs"]), K, N, sg.xc, _stream()))

    def acc():
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(prm), _ptr(sg.Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, sg.xc,
            _ptr(sg.ii_row), _ptr(sg.jj_row), _ptr(sg.idx), _ptr(sg.valid), _ptr(sg.Q), K, N,
            sg.E_local, 1, None, None, None, _ptr(sg.ws), sg.nbytes, _stream()))

    for _ in range(3):
        acc()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        acc()
    e1.record()
    torch.cuda.synchronize()
    ii = g["ii"][lo:hi].tolist()
    jj = g["jj"][lo:hi].tolist()
    d = idx.view(hi - lo, H, W)
    # spread of the gathered rows inside a 512-point tile (one image row): max - min row index
    rows = (d // W)
    spread = (rows.max(dim=2).values - rows.min(dim=2).values).float().mean().item()
    print(f"slice {r}: edges [{lo},{hi}) {e0.elapsed_time(e1) * 1e3 / 30:7.1f} us; distinct i {len(set(ii))}, "
          f"distinct j {len(set(jj))}; mean row spread of the matches of one image row {spread:.1f}; "
          f"valid {valid.float().mean().item():.2f}")
    del sg
