#!/usr/bin/env python
"""Timing of the tracker's pose solve on one B200, 512x384 frame:
  * frodo_track_pose (one persistent cooperative kernel), default stopping rule and a fixed
    50 iterations (rel_error = delta_norm = 0), rays and calib;
  * the same iteration written as the reference runs it -- PyTorch kernels on the GPU, one
    `.item()` per iteration (tracker.py:397-465; torch restatement below, because the
    reference's Python and lietorch do not travel to the GPU box).
CUDA events on the launch stream, 5 warm-up + 50 timed calls.  Prints one JSON line.
Usage: python profiles/bench_track.py [--out gpurun_out/track.json]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from frodo_slam_b200 import synth, tracking  # noqa: E402


def timed(fn, warm=5, n=50):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us


# ---- the reference's formulation in torch (tracker.py:397-465, geometry.py:17-52) ----------
def _act(T, X):
    q, w = T[3:6], T[6]
    uv = 2.0 * torch.linalg.cross(q.expand_as(X), X)
    return T[7] * (X + w * uv + torch.linalg.cross(q.expand_as(X), uv)) + T[0:3]


def _skew(x):
    o = torch.zeros_like(x[:, 0])
    return torch.stack([o, -x[:, 2], x[:, 1], x[:, 2], o, -x[:, 0], -x[:, 1], x[:, 0], o],
                       dim=-1).view(-1, 3, 3)


def _ray_dist(X, jac):
    d = torch.linalg.norm(X, dim=-1, keepdim=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = torch.cat((r, d), dim=-1)
    if not jac:
        return rd
    I = torch.eye(3, device=X.device).repeat(X.shape[0], 1, 1)
    dr = d_inv.unsqueeze(-1) * (I - (d_inv ** 2).unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
    return rd, torch.cat((dr, r.unsqueeze(-2)), dim=-2)


def torch_rays_loop(Xf, Xk, T, Qk, cfg, iters):
    """`iters` iterations of opt_pose_ray_dist_sim3 on compacted inputs; the retraction is
    replaced by a first-order update of t (it does not change the cost of the loop)."""
    sq = torch.sqrt(Qk)
    sqrt_info = torch.cat(((1 / cfg["sigma_ray"] * sq).repeat(1, 3), 1 / cfg["sigma_dist"] * sq), dim=1)
    rd_k = _ray_dist(Xk, False)
    T = T.clone()
    for _ in range(iters):
        P = _act(T, Xf)
        dP = torch.cat([torch.eye(3, device=P.device).repeat(P.shape[0], 1, 1), -_skew(P),
                        P.unsqueeze(-1)], dim=-1)
        rd_f, drd = _ray_dist(P, True)
        r = rd_k - rd_f
        J = -drd @ dP
        wr = sqrt_info * r
        a = wr.abs()
        hub = torch.where(a < cfg["huber"], torch.ones_like(a), cfg["huber"] / a)
        rsi = sqrt_info * torch.sqrt(hub)
        A = (rsi[..., None] * J).view(-1, 7)
        b = (rsi * r).view(-1, 1)
        H = A.T @ A
        g = -A.T @ b
        cost = 0.5 * (b.T @ b).item()  # host round trip, as in the reference
        L = torch.linalg.cholesky(H)
        tau = torch.cholesky_solve(g, L).view(-1)
        T[0:3] += 1e-3 * tau[0:3]
        _ = torch.linalg.norm(tau).item() < 0.0
    return T, cost


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    H, W = 384, 512
    N = H * W
    c = synth.tracking_case(H, W, seed=7, pose_noise=0.01)
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in c.items()}
    tp = dict(synth.TRACK_PARAMS)
    fixed = dict(tp, rel_error=0.0, delta_norm=0.0)
    Xf_c, Xk_c, meas, vmk = tracking.calib_measurements(t["Xf_all"], t["Xk"], t["K"], (H, W),
                                                        tp["depth_eps"])
    out = {"frame": "512x384", "points": N, "valid_fraction": float(c["valid"].float().mean())}

    def rays(cfg):
        return tracking.track_pose("rays", t["Xf_all"], t["Xk"], t["T_WCf"], t["T_WCk"], t["Qk"],
                                   t["valid"], idx_f2k=t["idx_f2k"], cfg=cfg)

    def calib(cfg):
        return tracking.track_pose("calib", Xf_c, Xk_c, t["T_WCf"], t["T_WCk"], t["Qk"], t["valid"],
                                   meas, vmk, t["K"], (H, W), idx_f2k=t["idx_f2k"], cfg=cfg)

    for name, fn in (("rays", rays), ("calib", calib)):
        it_default = int(fn(tp)["info"][0].item())
        us_default = timed(lambda: fn(tp))
        us_1 = timed(lambda: fn(dict(fixed, max_iters=1)))
        us_50 = timed(lambda: fn(fixed))
        per_iter = (us_50 - us_1) / 49.0
        # bytes the kernel must read once: Xf gather 12 + idx 8 + Xk|meas 12 + Q 4 + valid 1 (+1)
        alg = N * (12 + 8 + 12 + 4 + 1 + (1 if name == "calib" else 0))
        out[name] = {"default_rule": {"iterations": it_default, "us_per_solve": us_default},
                     "one_iteration_us": us_1, "fifty_iterations_us": us_50,
                     "us_per_iteration": per_iter,
                     "prologue_algorithmic_bytes": alg,
                     "on_chip_points_per_s": N / (per_iter * 1e-6)}
    # the reference's formulation (compacted inputs, as tracker.py:293-295)
    m = t["valid"][:, 0]
    Xf_v, Xk_v, Q_v = t["Xf_all"][t["idx_f2k"]][m].contiguous(), t["Xk"][m].contiguous(), t["Qk"][m].contiguous()
    T0 = t["T_WCf"].reshape(8)
    it = out["rays"]["default_rule"]["iterations"]
    us_t = timed(lambda: torch_rays_loop(Xf_v, Xk_v, T0, Q_v, tp, it), warm=2, n=10)
    us_t1 = timed(lambda: torch_rays_loop(Xf_v, Xk_v, T0, Q_v, tp, 1), warm=2, n=10)
    out["torch_formulation_rays"] = {"iterations": it, "us_per_solve": us_t, "one_iteration_us": us_t1,
                                     "note": "PyTorch kernels on the same GPU + one .item() per "
                                             "iteration, the way tracker.py:433-460 runs"}
    out["speedup_default_rule"] = us_t / out["rays"]["default_rule"]["us_per_solve"]
    line = json.dumps(out)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
