"""Records golden vectors from the UNMODIFIED reference backend (oracle/_ref, built by
oracle/build_ref.sh from /root/reference) on a CUDA device.  Run on the B200 box:

    python tests/golden/make_golden_ref.py gpurun_out/golden_ref.npz

and commit the result as tests/golden/golden_ref.npz.  The CPU tests then pin the oracle
(oracle/frodo_oracle.c) against these vectors.  Inputs are stored next to the outputs so the
fixture does not depend on the synthetic generator staying unchanged."""
import glob
import importlib.util
import io
import os
import sys
import contextlib

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import GP, MP, backend_gn, graph_inputs, matching_inputs  # noqa: E402


def load_ref():
    so = glob.glob(os.path.join(ROOT, "oracle", "_ref", "mast3r_slam_backends*.so"))[0]
    spec = importlib.util.spec_from_file_location("mast3r_slam_backends", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main(out_path):
    ref = load_ref()
    dev = torch.device("cuda:0")
    out = {}
    H, W = 24, 32  # N = 768, a multiple of the reference's 16-thread blocks (it has no bounds guard)
    d, rays, pts, p0 = matching_inputs(2, H, W, seed=21)
    p, c = ref.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), MP["max_iter"], MP["lambda_init"],
                         MP["convergence_thresh"])
    out.update(ip_rays=rays.numpy(), ip_pts=pts.numpy(), ip_p0=p0.numpy(), ip_p=p.cpu().numpy(),
               ip_conv=c.cpu().numpy())
    g = torch.Generator().manual_seed(5)
    p1 = (p.cpu().long() + torch.randint(-3, 4, p.shape, generator=g)).contiguous()
    D11 = d["D11"].half().contiguous()
    D21 = d["D21"].half().reshape(2, H * W, -1).contiguous()
    pn, sc = ref.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), MP["radius"],
                                MP["dilation_max"])
    out.update(rm_D11=D11.numpy(), rm_D21=D21.numpy(), rm_p1=p1.numpy(), rm_pn=pn.cpu().numpy(),
               rm_sc=sc.cpu().numpy())
    D11f, D21f = D11.float(), D21.float()
    pn32, sc32 = ref.refine_matches(D11f.to(dev), D21f.to(dev), p1.to(dev), 2, 3)
    out.update(rm32_pn=pn32.cpu().numpy(), rm32_sc=sc32.cpu().numpy())

    gr = graph_inputs("window8", H, W, seed=4, id_base=3, id_stride=2)
    for k in ("Twc", "Xs", "Cs", "K", "ii", "jj", "Q"):
        out["gn_" + k] = gr[k].numpy()
    out["gn_idx"] = gr["idx_ii2jj"].numpy().astype(np.int32)
    out["gn_valid"] = gr["valid_match"].numpy()
    out["gn_HW"] = np.array([H, W])
    sink = io.StringIO()
    for mode in ("rays", "points", "calib"):
        for iters in (1, 3):
            with contextlib.redirect_stdout(sink):
                Twc, dx = backend_gn(ref, mode, gr, dev, 1, iters)
            out[f"gn_{mode}_{iters}_Twc"] = Twc
            out[f"gn_{mode}_{iters}_dx"] = dx
    Twc, dx = backend_gn(ref, "rays", gr, dev, 2, 2)
    out["gn_rays_fix2_Twc"], out["gn_rays_fix2_dx"] = Twc, dx
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **out)
    print("wrote", out_path, os.path.getsize(out_path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden_ref.npz"))
