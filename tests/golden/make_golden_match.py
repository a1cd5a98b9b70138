#!/usr/bin/env python
"""Golden vectors for the matching pipeline around the two matching kernels
(mast3r_slam/matching.py:25-110: prep_for_iter_proj, the occlusion / threshold / linearise tail).

Runs the REFERENCE's own Python (imported from /root/reference, CPU torch) on small seeded inputs.
The reference's matching.py imports its CUDA extension at module import; this container has no
GPU, so the two kernels are served by the oracle (already pinned to the reference kernels'
outputs, tests/golden/golden_ref.npz) through a stub `mast3r_slam.backend` module.  Everything
else -- F.normalize, the Scharr gradient, index conversions, the distance test, the thresholds --
is the reference's code, unmodified.

Usage (in the build container only; /root/reference does not exist on the GPU box):
    python tests/golden/make_golden_match.py        -> tests/golden/golden_match.npz
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import oracle  # noqa: E402
from frodo_slam_b200 import synth  # noqa: E402

captured = {}


class _OracleBackends:
    """iter_proj / refine_matches with the pybind signatures (gn.cpp:120-150), on CPU."""

    @staticmethod
    def iter_proj(rays, pts, p_init, max_iter, lambda_init, cost_thresh):
        captured["rays_with_grad_img"] = rays.numpy().copy()
        captured["pts3d_norm"] = pts.numpy().copy()
        captured["p_init"] = p_init.numpy().copy()
        p, conv = oracle.iter_proj(rays.numpy(), pts.numpy(), p_init.numpy(), max_iter,
                                   lambda_init, cost_thresh)
        captured["p_iter"] = p.copy()
        captured["conv"] = conv.copy()
        return torch.from_numpy(p), torch.from_numpy(conv)

    @staticmethod
    def refine_matches(D11, D21, p1, radius, dilation_max):
        assert D11.dtype == torch.float16 and D21.dtype == torch.float16
        pn, sc = oracle.refine_matches(D11.numpy(), D21.numpy(), p1.numpy(), radius, dilation_max)
        captured["p_refined"] = pn.copy()
        captured["scores"] = sc.copy()
        return torch.from_numpy(pn), torch.from_numpy(sc)


stub = types.ModuleType("mast3r_slam.backend")
stub.mast3r_slam_backends = _OracleBackends
sys.modules["mast3r_slam.backend"] = stub

from mast3r_slam.config import load_config, config  # noqa: E402
cwd = os.getcwd()
os.chdir(REF)
load_config("config/base.yaml")
os.chdir(cwd)
import mast3r_slam.matching as ref_matching  # noqa: E402

out = {"cfg_" + k: np.float64(v) for k, v in config["matching"].items()}
B, H, W = 2, 32, 48
d = synth.matching_pairs(B, H, W, seed=3, device="cpu")
X11, X21 = d["X11"].float(), d["X21"].float()
D11, D21 = d["D11"].float(), d["D21"].float()
g = torch.Generator().manual_seed(5)
ident = torch.arange(H * W)[None].repeat(B, 1)
jitter = torch.randint(-2, 3, (B, H * W), generator=g) + W * torch.randint(-2, 3, (B, H * W), generator=g)
idx_init = (ident + jitter).clamp(0, H * W - 1)

for tag, init in (("a", None), ("b", idx_init)):
    captured.clear()
    idx, valid = ref_matching.match(X11, X21, D11, D21, init)
    out[f"{tag}_idx_1_to_2"] = idx.numpy()
    out[f"{tag}_valid"] = valid.numpy()
    for k, v in captured.items():
        out[f"{tag}_{k}"] = v
out.update(X11=X11.numpy(), X21=X21.numpy(), D11=D11.numpy().astype(np.float16),
           D21=D21.numpy().astype(np.float16), idx_init=idx_init.numpy())
# descriptors are stored already rounded to fp16 (what D.half() yields): halves the file
path = os.path.join(ROOT, "tests", "golden", "golden_match.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes;",
      {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if k.startswith("a_")})
print("valid fraction", out["a_valid"].mean(), out["b_valid"].mean())
