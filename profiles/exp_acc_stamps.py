#!/usr/bin/env python
"""Per-CTA start / end-of-loop / end stamps (globaltimer) of the accumulate kernel, from a
library built with -DFRODO_ACC_STAMPS (profiles/experiments/bin/libfrodo_stamps.so)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from frodo_slam_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "profiles", "experiments", "bin", os.environ.get("STAMPLIB", "libfrodo_stamps.so"))
from frodo_slam_b200 import backends, dist as fdist, synth
from frodo_slam_b200.backends import _ptr, _stream
dev = torch.device("cuda:0")
H, W = 384, 512
gp = synth.GN_PARAMS
lib = _lib.load()
raw = ctypes.CDLL(_lib.LIB_PATH)
_keep = []
_sm_means = []
for cfg in os.environ.get("CFGS", "window8").split(","):
    if cfg.startswith("shift:"):
        assert raw.frodo_debug_acc_shift(int(cfg[6:])) == 0; continue
    if cfg.startswith("pad:"):
        _keep.append(torch.empty(int(cfg[4:]) << 20, dtype=torch.uint8, device=dev)); continue
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
    for _ in range(5): acc()
    torch.cuda.synchronize()
    ncta = min(16384, int(os.environ.get("NCTA", "16384")))
    buf = np.zeros(4 * 16384, dtype=np.uint64)
    rc = raw.frodo_debug_acc_stamps(buf.ctypes.data_as(ctypes.c_void_p), 16384)
    assert rc == 0
    st = buf.reshape(-1, 4)
    st = st[st[:, 0] > 0]
    t0 = st[:, 0].min()
    start = (st[:, 0] - t0).astype(np.float64) / 1e3
    loop_end = (st[:, 1] - t0).astype(np.float64) / 1e3
    end = (st[:, 2] - t0).astype(np.float64) / 1e3
    sm = st[:, 3].astype(np.int64)
    print(f"{cfg}: E={E} CTAs={len(st)} ppc={os.environ.get('FRODO_GN_PPC','auto')}")
    print(f"  start  us: min {start.min():.1f} p50 {np.median(start):.1f} p90 {np.percentile(start,90):.1f} max {start.max():.1f}")
    dur = loop_end - start
    print(f"  loop   us: min {dur.min():.1f} p50 {np.median(dur):.1f} p90 {np.percentile(dur,90):.1f} max {dur.max():.1f}")
    ep = end - loop_end
    print(f"  epilog us: min {ep.min():.1f} p50 {np.median(ep):.1f} p90 {np.percentile(ep,90):.1f} max {ep.max():.1f}")
    print(f"  end    us: p50 {np.median(end):.1f} max {end.max():.1f}")
    per_sm = np.bincount(sm, minlength=148)
    print(f"  CTAs per SM: min {per_sm.min()} max {per_sm.max()}  SMs used {np.count_nonzero(per_sm)}")
    # first-wave CTAs (start < 5 us) vs later ones
    first = start < 5.0
    print(f"  first-wave CTAs {first.sum()}  their loop p50 {np.median(dur[first]):.1f}; "
          f"later CTAs {np.count_nonzero(~first)} loop p50 "
          f"{np.median(dur[~first]) if (~first).any() else float('nan'):.1f}")
    # per-SM finishing time
    sm_end = np.zeros(148); 
    for s_, e_ in zip(sm, end): sm_end[s_] = max(sm_end[s_], e_)
    print(f"  per-SM last end: min {sm_end[sm_end>0].min():.1f} p50 {np.median(sm_end[sm_end>0]):.1f} max {sm_end.max():.1f}")
    if os.environ.get("DECOMP"):
        nch = len(st) // E
        full = buf.reshape(-1, 4)[:E * nch]
        d = ((full[:, 1] - full[:, 0]).astype(np.float64) / 1e3).reshape(E, nch)
        smm = full[:, 3].astype(np.int64).reshape(E, nch)
        tot = d.var()
        print(f"  variance total {tot:.1f}; explained by edge {d.mean(1).var():.1f}, by chunk {d.mean(0).var():.1f}")
        sm_mean = np.array([d[smm == s_].mean() if (smm == s_).any() else np.nan for s_ in range(148)])
        print(f"  explained by SM {np.nanvar(sm_mean):.1f}")
        _sm_means.append(sm_mean)
        if len(_sm_means) > 1:
            a, b = _sm_means[0], sm_mean
            ok = ~np.isnan(a) & ~np.isnan(b)
            print(f"  correlation of SM means with first run: {np.corrcoef(a[ok], b[ok])[0,1]:.3f}")
        print("  edge means:", " ".join(f"{x:.0f}" for x in d.mean(1)))
        print("  chunk means:", " ".join(f"{x:.0f}" for x in d.mean(0)))
        print("  SM means:", " ".join(f"{x:.0f}" for x in sm_mean))
        print("  ii:", g["ii"].tolist()); print("  jj:", g["jj"].tolist())
    del sg, g
