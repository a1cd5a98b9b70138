import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from frodo_slam_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "profiles/experiments/bin/libfrodo_ss.so")
from frodo_slam_b200 import backends, synth
dev = torch.device("cuda:0"); gp = synth.GN_PARAMS
raw = ctypes.CDLL(_lib.LIB_PATH)
for cfg in ("window8", "pair"):
    K, und = synth.graph_edges(cfg)
    g = synth.keyframe_graph(K, 32, 32, und, seed=1, device=dev)
    for _ in range(3):
        T = g["Twc"].clone()
        backends.gauss_newton_rays(T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], gp["sigma_ray"], gp["sigma_dist"], gp["C_conf"], gp["Q_conf"], 1, 10, 0.0)
    torch.cuda.synchronize()
    st = np.zeros(16, dtype=np.int64)
    assert raw.frodo_debug_small_stamps(st.ctypes.data_as(ctypes.c_void_p)) == 0
    names = ["stage", "build", "factor0", "products", "factor1", "backsub", "finish"]
    print(cfg, " ".join(f"{n} {(st[i+1]-st[i])/1965.0:.2f}" for i, n in enumerate(names)), "total", (st[7]-st[0])/1965.0)
    print("   build detail (us from stage end): " + " ".join(f"tile{t}: lists {(st[8+t]-st[1])/1965.0:.2f} done {(st[11+t]-st[1])/1965.0:.2f}" for t in range(3)))
