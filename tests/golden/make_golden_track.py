#!/usr/bin/env python
"""Golden vectors for the tracker's frame-to-keyframe pose solve:
FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 / solve (mast3r_slam/tracker.py:397-517)
with geometry.py (point_to_ray_dist, act_Sim3, project_calib, constrain_points_to_ray) and
nonlinear_optimizer.py (check_convergence, huber).

Runs the REFERENCE's own Python, imported from /root/reference, on CPU torch.  Two things the
reference needs are absent from this container and are stubbed:
  * `lietorch` (unpinned git dependency, pyproject.toml:15): a minimal torch `Sim3` with the four
    group operations the tracker uses (inv, *, act, retr) written from the published formulas --
    the same ones gn_kernels.cu:219-314 (quat_comp, actSO3, actSim3, relSim3) and :341-455
    (expSim3, retrSim3) carry.  Exp / retr go through the oracle's pose_retr, which is pinned
    to the reference's pose_retr_kernel by golden_ref.npz.
  * the modules tracker.py imports but the two methods never touch (frame containers, the MASt3R
    network, the viewer, matplotlib, loguru ...): MagicMock placeholders.
Everything that forms residuals, Jacobians, weights, the normal equations, the Cholesky solve and
the convergence test is the reference's code, unmodified.

Usage (build container only; /root/reference does not exist on the GPU box):
    python tests/golden/make_golden_track.py        -> tests/golden/golden_track.npz
"""
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import oracle  # noqa: E402
from frodo_slam_b200 import synth  # noqa: E402


def _act(T, X):
    q = T[..., 3:6]
    w = T[..., 6:7]
    uv = 2.0 * torch.cross(q.expand_as(X), X, dim=-1)
    Y = X + w * uv + torch.cross(q.expand_as(X), uv, dim=-1)
    return T[..., 7:8] * Y + T[..., 0:3]


def _qmul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], dim=-1)


class Sim3:
    """Stand-in for lietorch.Sim3: data [...,8] = [t, q_xyzw, s]."""

    def __init__(self, data):
        self.data = data

    def inv(self):
        d = self.data
        qi = d[..., 3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0])
        si = 1.0 / d[..., 7:8]
        rot_only = torch.cat([torch.zeros_like(d[..., 0:3]), qi, torch.ones_like(si)], dim=-1)
        t = -si * _act(rot_only, d[..., 0:3])
        return Sim3(torch.cat([t, qi, si], dim=-1))

    def __mul__(self, other):
        a, b = self.data, other.data
        return Sim3(torch.cat([_act(a, b[..., 0:3]), _qmul(a[..., 3:7], b[..., 3:7]),
                               a[..., 7:8] * b[..., 7:8]], dim=-1))

    def act(self, p):
        return _act(self.data, p)

    def retr(self, tau):
        out = oracle.pose_retr(self.data.reshape(-1, 8).numpy(), tau.reshape(-1, 7).numpy(), 0)
        return Sim3(torch.from_numpy(out).reshape(self.data.shape))


lie = types.ModuleType("lietorch")
lie.Sim3 = Sim3
sys.modules["lietorch"] = lie
for name in ("mast3r_slam.frame", "mast3r_slam.local_mapping", "mast3r_slam.mast3r_utils", "mast3r",
             "mast3r.fast_nn", "mast3r_slam.visualization_utils", "matplotlib", "matplotlib.pyplot",
             "mast3r_slam.matching", "mast3r_slam.pgo", "loguru", "mast3r_slam.height_prior"):
    sys.modules[name] = MagicMock()

from mast3r_slam.config import load_config, config  # noqa: E402
cwd = os.getcwd()
os.chdir(REF)
load_config("config/base.yaml")
os.chdir(cwd)
import mast3r_slam.tracker as ref_tracker  # noqa: E402
from mast3r_slam.geometry import constrain_points_to_ray, get_pixel_coords  # noqa: E402

FT = ref_tracker.FrameTracker


class _Self:
    """Carries what the two methods read from `self`: cfg and solve()."""
    cfg = config["tracking"]

    def __init__(self):
        self.costs = []

    def solve(self, sqrt_info, r, J):
        tau, cost = FT.solve(self, sqrt_info, r, J)
        self.costs.append(cost)
        return tau, cost


def run_case(tag, H, W, seed, pose_noise, out):
    c = synth.tracking_case(H, W, seed=seed, pose_noise=pose_noise)
    valid = c["valid"]
    m = valid[:, 0]
    T_WCf, T_WCk = Sim3(c["T_WCf"].clone()), Sim3(c["T_WCk"].clone())
    # --- uncalibrated: tracker.py:291-298 hands over the compacted arrays
    s = _Self()
    with torch.inference_mode():
        Tn, Trel = FT.opt_pose_ray_dist_sim3(s, c["Xf"][m], c["Xk"][m], T_WCf, T_WCk,
                                             c["Qk"][m], valid)
    out[f"{tag}_rays_T_WCf"] = Tn.data.numpy().copy()
    out[f"{tag}_rays_T_CkCf"] = Trel.data.numpy().copy()
    out[f"{tag}_rays_costs"] = np.array(s.costs, np.float64)
    # --- calibrated: the branch of get_points_poses, tracker.py:383-395, then :300-311
    Kmat = c["K"]
    img_size = (H, W)
    with torch.inference_mode():
        Xf_c = constrain_points_to_ray(img_size, c["Xf_all"][None], Kmat).squeeze(0)
        Xk_c = constrain_points_to_ray(img_size, c["Xk"][None], Kmat).squeeze(0)
        uv_k = get_pixel_coords(1, img_size, device=Xf_c.device, dtype=Xf_c.dtype).view(-1, 2)
        meas_k = torch.cat((uv_k, torch.log(Xk_c[..., 2:3])), dim=-1)
        valid_meas_k = Xk_c[..., 2:3] > config["tracking"]["depth_eps"]
        meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
        s = _Self()
        Tn, Trel = FT.opt_pose_calib_sim3(s, Xf_c[c["idx_f2k"]], Xk_c, T_WCf, T_WCk, c["Qk"],
                                          valid, meas_k, valid_meas_k, Kmat, img_size)
    out[f"{tag}_calib_T_WCf"] = Tn.data.numpy().copy()
    out[f"{tag}_calib_T_CkCf"] = Trel.data.numpy().copy()
    out[f"{tag}_calib_costs"] = np.array(s.costs, np.float64)
    out[f"{tag}_calib_Xf"] = Xf_c.numpy().copy()
    out[f"{tag}_calib_Xk"] = Xk_c.numpy().copy()
    out[f"{tag}_calib_meas_k"] = meas_k.numpy().copy()
    out[f"{tag}_calib_valid_meas_k"] = valid_meas_k.numpy().copy()
    for k in ("Xf_all", "idx_f2k", "Xk", "Qk", "valid", "T_WCf", "T_WCk", "K"):
        out[f"{tag}_{k}"] = c[k].numpy().copy()
    out[f"{tag}_HW"] = np.array([H, W])


out = {"cfg_" + k: np.float64(v) for k, v in config["tracking"].items()
       if isinstance(v, (int, float)) and not isinstance(v, bool)}
run_case("a", 32, 48, seed=1, pose_noise=0.01, out=out)
run_case("b", 40, 56, seed=4, pose_noise=0.03, out=out)
path = os.path.join(ROOT, "tests", "golden", "golden_track.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes")
for t in "ab":
    print(t, "rays iters", len(out[f"{t}_rays_costs"]), out[f"{t}_rays_costs"],
          "calib iters", len(out[f"{t}_calib_costs"]), out[f"{t}_calib_costs"])
