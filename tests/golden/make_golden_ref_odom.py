"""Records golden vectors of gauss_newton_rays_odom from the UNMODIFIED reference backend
(oracle/_ref) on a CUDA device, like make_golden_ref.py:

    python tests/golden/make_golden_ref_odom.py gpurun_out/golden_ref_odom.npz

Commit the result as tests/golden/golden_ref_odom.npz."""
import contextlib
import io
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

from helpers import backend_gn_odom, graph_inputs, odom_inputs  # noqa: E402
from make_golden_ref import load_ref  # noqa: E402


def main(out_path):
    ref = load_ref()
    dev = torch.device("cuda:0")
    H, W = 24, 32
    g = graph_inputs("window8", H, W, seed=6)
    od = odom_inputs(g, seed=9)
    out = {}
    for k in ("Twc", "Xs", "Cs", "ii", "jj", "Q"):
        out["gn_" + k] = g[k].numpy()
    out["gn_idx"] = g["idx_ii2jj"].numpy().astype(np.int32)
    out["gn_valid"] = g["valid_match"].numpy()
    for k in ("odom_ii", "odom_jj", "Tij", "s_bar"):
        out[k] = od[k].cpu().numpy()
    sink = io.StringIO()
    # the reference prints from C++ (std::cout): silence fd 1 around the calls
    sys.stdout.flush()
    saved = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)
    os.dup2(null, 1)
    try:
        for num_fix, iters in ((1, 1), (1, 3), (2, 2)):
            with contextlib.redirect_stdout(sink):
                Twc, dx = backend_gn_odom(ref, g, od, dev, num_fix, iters)
            out[f"odom_fix{num_fix}_it{iters}_Twc"] = Twc
            out[f"odom_fix{num_fix}_it{iters}_dx"] = dx
    finally:
        os.dup2(saved, 1)
        os.close(null)
        os.close(saved)
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else
         os.path.join(ROOT, "gpurun_out", "golden_ref_odom.npz"))
