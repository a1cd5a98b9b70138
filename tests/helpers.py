"""Shared input builders for the parity tests (seeded, small)."""
import numpy as np
import torch

from frodo_slam_b200 import synth

MP = synth.MATCH_PARAMS
GP = synth.GN_PARAMS


def matching_inputs(B, H, W, seed=0, init="identity"):
    d = synth.matching_pairs(B, H, W, seed=seed)
    idx_init = None
    if init == "random":
        g = torch.Generator().manual_seed(seed + 100)
        idx_init = torch.randint(0, H * W, (B, H * W), generator=g)
    rays, pts, p0 = synth.prep_for_iter_proj(d["X11"], d["X21"], idx_init)
    if init == "outside":  # exercise the clamp of the initial guess
        p0 = p0 * 1.5 - 10.0
    return d, rays, pts, p0.contiguous()


def graph_inputs(cfg, H, W, seed=1, **kw):
    K, und = synth.graph_edges(cfg)
    return synth.keyframe_graph(K, H, W, und, seed=seed, **kw)


def gn_args(mode, g):
    """(oracle kwargs, backend positional tail) for the three GN ops."""
    if mode == "rays":
        return dict(sigma_a=GP["sigma_ray"], sigma_b=GP["sigma_dist"]), None
    if mode == "points":
        return dict(sigma_a=GP["sigma_point"], sigma_b=0.0), None
    return dict(sigma_a=GP["sigma_pixel"], sigma_b=GP["sigma_depth"], K=g["K"].numpy(),
                height=g["H"], width=g["W"], pixel_border=GP["pixel_border"],
                z_eps=GP["depth_eps"]), None


def oracle_gn(mode, g, num_fix, max_iter, delta_thresh=0.0, return_system=False):
    from oracle import oracle
    kw, _ = gn_args(mode, g)
    return oracle.gauss_newton(mode, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(),
                               g["ii"].numpy(), g["jj"].numpy(), g["idx_ii2jj"].numpy(),
                               g["valid_match"].numpy(), g["Q"].numpy(), C_thresh=GP["C_conf"],
                               Q_thresh=GP["Q_conf"], num_fix=num_fix, max_iter=max_iter,
                               delta_thresh=delta_thresh, return_system=return_system, **kw)


def backend_gn(be, mode, g, dev, num_fix, max_iter, delta_thresh=0.0):
    """Calls the operator exactly as the reference callers do; returns (Twc, dx) numpy."""
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    Twc = t["Twc"].clone()
    common = (Twc, t["Xs"], t["Cs"])
    edges = (t["ii"], t["jj"], t["idx_ii2jj"], t["valid_match"], t["Q"])
    if mode == "rays":
        dx = be.gauss_newton_rays(*common, *edges, GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"],
                                  GP["Q_conf"], num_fix, max_iter, delta_thresh)
    elif mode == "points":
        assert num_fix == 1
        dx = be.gauss_newton_points(*common, *edges, GP["sigma_point"], GP["C_conf"], GP["Q_conf"],
                                    max_iter, delta_thresh)
    else:
        assert num_fix == 1
        dx = be.gauss_newton_calib(*common, t["K"], *edges, g["H"], g["W"], GP["pixel_border"],
                                   GP["depth_eps"], GP["sigma_pixel"], GP["sigma_depth"],
                                   GP["C_conf"], GP["Q_conf"], max_iter, delta_thresh)
    torch.cuda.synchronize()
    return Twc.cpu().numpy(), dx[0].cpu().numpy()


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def expand_edge_blocks(eb):
    """[E,35] (28 upper-tri of A, 7 of g_j) -> Hs [4,E,7,7], gs [2,E,7] in the reference layout."""
    eb = np.asarray(eb)
    E = eb.shape[0]
    A = np.zeros((E, 7, 7))
    iu = np.triu_indices(7)
    A[:, iu[0], iu[1]] = eb[:, :28]
    A[:, iu[1], iu[0]] = eb[:, :28]
    gj = eb[:, 28:]
    return np.stack([A, -A, -A, A]), np.stack([-gj, gj])


OP = synth.ODOM_PARAMS


def odom_inputs(g, **kw):
    return synth.odometry_edges(g, **kw)


def oracle_gn_odom(g, od, num_fix, max_iter, delta_thresh=0.0):
    from oracle import oracle
    return oracle.gauss_newton_rays_odom(
        g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), g["ii"].numpy(), g["jj"].numpy(),
        g["idx_ii2jj"].numpy(), g["valid_match"].numpy(), g["Q"].numpy(), od["odom_ii"].numpy(),
        od["odom_jj"].numpy(), od["Tij"].numpy(), od["s_bar"].numpy(), OP["sigma_odom_t"],
        OP["sigma_odom_r"], OP["sigma_ray"], GP["sigma_dist"], OP["sigma_scale_prior"],
        GP["C_conf"], GP["Q_conf"], num_fix, max_iter, delta_thresh)


def backend_gn_odom(be, g, od, dev, num_fix, max_iter, delta_thresh=0.0):
    """gauss_newton_rays_odom called as global_opt.py:260-283 does (s_bar stays a CPU tensor)."""
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    Twc = t["Twc"].clone()
    dx = be.gauss_newton_rays_odom(
        Twc, t["Xs"], t["Cs"], t["ii"], t["jj"], t["idx_ii2jj"], t["valid_match"], t["Q"],
        od["odom_ii"].to(dev), od["odom_jj"].to(dev), od["Tij"].to(dev), od["s_bar"],
        OP["sigma_odom_t"], OP["sigma_odom_r"], OP["sigma_ray"], GP["sigma_dist"],
        OP["sigma_scale_prior"], GP["C_conf"], GP["Q_conf"], num_fix, max_iter, delta_thresh)
    torch.cuda.synchronize()
    return Twc.cpu().numpy(), dx[0].cpu().numpy()
