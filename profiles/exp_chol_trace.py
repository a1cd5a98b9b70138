#!/usr/bin/env python
"""Event trace of the dataflow Cholesky (library built with -DFRODO_CHOL_TRACE): follows, for a
column k, the chain W_k -> L(k+2,k) -> update of (k+2,k+1) -> hand-over -> spine, with
globaltimer stamps of the CTAs involved (last launch of a 10-iteration solve, 512x384 loop64)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from frodo_slam_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "profiles", "experiments", "bin", "libfrodo_choltrace.so")
from frodo_slam_b200 import backends, synth
from frodo_slam_b200.backends import _ptr, _stream
dev = torch.device("cuda:0")
H, W = 384, 512
gp = synth.GN_PARAMS
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH)
K, und = synth.graph_edges("loop64")
g = synth.keyframe_graph(K, H, W, und, seed=1, device=dev)
for _ in range(2):
    T_ = g["Twc"].clone()
    backends.gauss_newton_rays(T_, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                               g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"],
                               gp["C_conf"], gp["Q_conf"], 1, 10, 0.0)
torch.cuda.synchronize()
buf = np.zeros(160 * 128 * 2, dtype=np.uint64)
cnt = np.zeros(160, dtype=np.int32)
assert raw.frodo_debug_chol_trace(buf.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p)) == 0
tr = buf.reshape(160, 128, 2)
ev = {}      # (cta) -> list of (kind, m, aux, t)
owner = {}   # (i, j) -> cta
for c in range(160):
    L = []
    for e in range(min(cnt[c], 128)):
        code, t = int(tr[c, e, 0]), int(tr[c, e, 1])
        L.append((code >> 16, (code >> 8) & 255, code & 255, t))
    ev[c] = L
    if c not in (0, 159) and L and L[0][0] == 1:
        owner[(L[0][1], L[0][2])] = c
def find(c, kind, m, aux=None):
    for (k_, m_, a_, t) in ev[c]:
        if k_ == kind and m_ == m and (aux is None or a_ == aux):
            return t
    return None
print(f"CTAs with events: {sum(1 for c in ev if ev[c])}; tiles owned: {len(owner)}")
names = {}
for k in [int(x) for x in os.environ.get("COLS", "4,5,6,7,8").split(",")]:
    t0 = find(0, 21, k)
    rel = lambda t: "   n/a" if t is None else f"{(t - t0) / 1e3:6.2f}"
    print(f"--- column {k}: times (us) relative to 'W_{k} published' by the spine")
    print(f"  spine: factor done {rel(find(0, 20, k))}  W pub {rel(t0)}  L({k+1},{k}) pub {rel(find(0, 22, k))}  "
          f"diag upd {rel(find(0, 23, k))}  next factor done {rel(find(0, 20, k + 1))}  next W pub {rel(find(0, 21, k + 1))}")
    print(f"  spine fetch warp: sub-diag ({k+2},{k+1}) seen {rel(find(159, 24, k + 1))}  diag ({k+2},{k+2}) seen {rel(find(159, 25, k + 1))}")
    a = owner.get((k + 2, k))
    if a is not None:
        print(f"  worker of ({k+2},{k}) [cta {a}]: wait W from {rel(find(a, 4, k))}  flag seen {rel(find(a, 5, k))}  "
              f"loaded {rel(find(a, 6, k))}  L({k+2},{k}) pub {rel(find(a, 7, k, k + 2))}")
    b = owner.get((k + 2, k + 1))
    if b is not None:
        print(f"  worker of ({k+2},{k+1}) [cta {b}]: wait L(.,{k}) from {rel(find(b, 8, k))}  seen {rel(find(b, 9, k))}  "
              f"loaded {rel(find(b, 10, k))}  applied {rel(find(b, 11, k))}  hand-over pub {rel(find(b, 3, k + 1, 1))}")
    c = owner.get((k + 2, k + 2))
    if c is not None:
        print(f"  worker of ({k+2},{k+2}) [cta {c}]: wait L(.,{k}) from {rel(find(c, 8, k))}  seen {rel(find(c, 9, k))}  "
              f"loaded {rel(find(c, 10, k))}  applied {rel(find(c, 11, k))}  hand-over pub {rel(find(c, 3, k + 2, 0))}")
