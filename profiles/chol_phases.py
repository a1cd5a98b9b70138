#!/usr/bin/env python
"""Phase profile of the dataflow Cholesky (chol_flow_kernel, csrc/gn_chol_flow.cu): reads the
clock64 stamps the spine CTA leaves in the workspace (FRODO_CHOL_STAMPS=1) after one GN
iteration and prints the time per phase: per tile column the diagonal factor + inversion, the
sub-diagonal panel product + publication, the update of the next diagonal tile; then the back
substitution and the pose update.  Also times the kernel with CUDA events.
Usage: FRODO_CHOL_STAMPS=1 python profiles/chol_phases.py [--kf loop64] [--hw 96x128]"""
import argparse
import ctypes
import os
import sys

os.environ.setdefault("FRODO_CHOL_STAMPS", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--kf", default="loop64")
ap.add_argument("--hw", default="96x128")
ap.add_argument("--insitu", action="store_true",
                help="stamps of the LAST iteration of a whole 10-iteration frodo_gauss_newton call "
                     "(tiles built from the per-edge blocks, caches as the accumulate leaves them)")
args = ap.parse_args()
H, W = (int(x) for x in args.hw.split("x"))

from frodo_slam_b200 import _lib, backends, synth  # noqa: E402
from frodo_slam_b200.backends import _ptr, _stream  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda:0")
gp = synth.GN_PARAMS
K, und = synth.graph_edges(args.kf)
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
E, N = g["ii"].shape[0], H * W
prm = backends.make_params("rays", sigma_ray=gp["sigma_ray"], sigma_dist=gp["sigma_dist"],
                           C_thresh=gp["C_conf"], Q_thresh=gp["Q_conf"])
n = 7 * (K - 1)
T, Tc = (n + 1 + 31) // 32, (n + 31) // 32
nbytes = lib.frodo_gn_workspace_bytes(K, E, N)
ws = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
st_off = lib.frodo_gn_chol_stamps_offset(K)
Hg = torch.empty((n + 1) * n, dtype=torch.float64, device=dev)
eb = torch.empty((E, 35), dtype=torch.float64, device=dev)
ii_row = torch.empty(E, dtype=torch.int32, device=dev)
jj_row = torch.empty(E, dtype=torch.int32, device=dev)
dx = torch.zeros(K - 1, 7, device=dev)
s = _stream()
_lib.check(lib.frodo_gn_workspace_init(_ptr(ws), nbytes, K, E, N, s))
_lib.check(lib.frodo_gn_edge_index(_ptr(g["ii"]), _ptr(g["jj"]), E, _ptr(ii_row), _ptr(jj_row), None, s))
Twc = g["Twc"].clone()


def system():
    _lib.check(lib.frodo_gn_accumulate(
        ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, None, _ptr(ii_row),
        _ptr(jj_row), _ptr(g["idx_ii2jj"]), _ptr(g["valid_match"]), _ptr(g["Q"]), K, N, E, 1,
        _ptr(Hg), _ptr(eb), None, _ptr(ws), nbytes, s))


def solve():
    _lib.check(lib.frodo_gn_solve_retract(_ptr(Hg), _ptr(Twc), _ptr(dx), K, 1, 0.0, None, None,
                                          _ptr(ws), nbytes, s))


times = []
if args.insitu:
    for rep in range(3):
        Twc.copy_(g["Twc"])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.frodo_gauss_newton(
            ctypes.byref(prm), _ptr(Twc), _ptr(g["Xs"]), _ptr(g["Cs"]), None, _ptr(g["ii"]),
            _ptr(g["jj"]), _ptr(g["idx_ii2jj"]), _ptr(g["valid_match"]), _ptr(g["Q"]), K, N, E, 1,
            10, 0.0, _ptr(dx), None, _ptr(ws), nbytes, s))
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e3 / 10)
    print("IN SITU: whole call / 10 iterations (us per iteration) follows as 'kernel' times")
for it in range(0 if args.insitu else 6):
    Twc.copy_(g["Twc"])
    system()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    solve()
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1) * 1e3)
st = ws[st_off:st_off + 1024 * 8].view(torch.int64).cpu().numpy().astype("float64")
mhz = 1965.0
us = lambda a, b: (b - a) / mhz  # noqa: E731
print(f"n={n} T={T} Tc={Tc}; kernel by CUDA events (us): " + " ".join(f"{t:.1f}" for t in times))
print(f"(clock64 cycles of the spine CTA at {mhz:.0f} MHz -> us; "
      + ("tiles from per-edge blocks, in situ)" if args.insitu else "dense-Hg tile build)"))
print(f"wait for + load the first diagonal tile {us(st[0], st[1]):8.2f} us")
tot = {"factor+invert": 0.0, "panel+publish": 0.0, "diag update": 0.0}
for k in range(Tc):
    a, b, c, d = (st[2 + 4 * k + i] for i in range(4))
    c = c if c > 0 else b
    tot["factor+invert"] += us(a, b)
    tot["panel+publish"] += us(b, c)
    tot["diag update"] += us(c, d)
    print(f"col {k:2d}: factor+invert (+fetch) {us(a, b):6.2f}  panel+publish {us(b, c):6.2f}  "
          f"diag update {us(c, d):6.2f}")
print({k: round(float(v), 2) for k, v in tot.items()})
print("warp 0: factor start -> end (us) / start-to-start period:")
print(" ".join(f"{us(st[768 + 2 * k], st[768 + 2 * k + 1]):.2f}/{us(st[768 + 2 * k], st[768 + 2 * k + 2]) if k + 1 < Tc else 0:.2f}"
               for k in range(Tc)))
f_end, b_end, end = st[2 + 4 * Tc], st[3 + 4 * Tc], st[4 + 4 * Tc]
print(f"factorisation total {us(st[0], f_end):8.2f} us")
print(f"back substitution   {us(f_end, b_end):8.2f} us  (flag sweep {us(f_end, st[100]):.2f}; columns "
      + " ".join(f"{us(st[101 + k + 1] if k + 1 < Tc else st[100], st[101 + k]):.2f}" for k in range(Tc - 1, -1, -1)) + ")")
print(f"pose update         {us(b_end, end):8.2f} us")
print(f"spine total         {us(st[0], end):8.2f} us")
