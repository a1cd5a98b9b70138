"""Seeded synthetic inputs for the matching and keyframe-graph ops.

Scene: cameras sit near the centre of a sphere (radius 4, centre (0,0,-1)) and
look along +z, so every pixel ray hits the surface at a closed-form depth and
the true correspondence between any two keyframes is known analytically.
Descriptors are a smooth 24-channel field attached to the surface.

All random draws come from a CPU ``torch.Generator`` so the same seed gives the
same inputs on every machine; the heavy per-pixel math runs on ``device``.
"""
import math

import torch

SPHERE_C = (0.0, 0.0, -1.0)
SPHERE_R = 4.0

# config/base.yaml:9-16 and :40-56 of the reference
MATCH_PARAMS = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=0.1,
                    radius=3, dilation_max=5, ft_matching_threshold=0.8)
GN_PARAMS = dict(sigma_ray=0.003, sigma_dist=10.0, sigma_point=0.05, sigma_pixel=1.0,
                 sigma_depth=10.0, C_conf=0.0, Q_conf=1.5, pixel_border=-10, depth_eps=1e-6,
                 max_iters=10, delta_norm=1e-8, num_fix=1)


def tum_intrinsics(H, W, dtype=torch.float32):
    """config/intrinsics.yaml: [517.3,516.5,318.6,255.3] @ 640x480, scaled to HxW."""
    sx, sy = W / 640.0, H / 480.0
    K = torch.tensor([[517.3 * sx, 0.0, 318.6 * sx], [0.0, 516.5 * sy, 255.3 * sy],
                      [0.0, 0.0, 1.0]], dtype=dtype)
    return K


def quat_to_R(q):
    """q = [x,y,z,w] -> 3x3 (any leading dims)."""
    x, y, z, w = q.unbind(-1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], dim=-1)
    return R.reshape(q.shape[:-1] + (3, 3))


def random_poses(n, gen, rot_deg=3.0, trans=0.05, scale=0.05):
    """[n,8] Sim3 poses [t(3), q_xyzw(4), s] (float64), pose 0 close to identity too."""
    axis = torch.randn(n, 3, generator=gen, dtype=torch.float64)
    axis = axis / axis.norm(dim=-1, keepdim=True)
    ang = (torch.rand(n, 1, generator=gen, dtype=torch.float64) * 2 - 1) * math.radians(rot_deg)
    q = torch.cat([axis * torch.sin(ang / 2), torch.cos(ang / 2)], dim=-1)
    t = (torch.rand(n, 3, generator=gen, dtype=torch.float64) * 2 - 1) * trans
    s = 1.0 + (torch.rand(n, 1, generator=gen, dtype=torch.float64) * 2 - 1) * scale
    return torch.cat([t, q, s], dim=-1)


def pixel_rays(K, H, W, device):
    v, u = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float64),
                          torch.arange(W, device=device, dtype=torch.float64), indexing="ij")
    K = K.to(device=device, dtype=torch.float64)
    x = (u - K[0, 2]) / K[0, 0]
    y = (v - K[1, 2]) / K[1, 1]
    return torch.stack([x, y, torch.ones_like(x)], dim=-1).reshape(-1, 3)  # [N,3]


def pointmap(K, T, H, W, device):
    """Camera-frame points X_c [N,3] and world points X_w [N,3] (float64) for pose T [8]."""
    T = T.to(device=device, dtype=torch.float64)
    d = pixel_rays(K, H, W, device)
    R = quat_to_R(T[3:7])
    t, s = T[0:3], T[7]
    c = torch.tensor(SPHERE_C, device=device, dtype=torch.float64)
    Rd = d @ R.T
    a = (s * s) * (d * d).sum(-1)
    b = 2 * s * (Rd * (t - c)).sum(-1)
    cc = ((t - c) ** 2).sum() - SPHERE_R ** 2
    lam = (-b + torch.sqrt(b * b - 4 * a * cc)) / (2 * a)
    Xc = lam[:, None] * d
    Xw = s * (Xc @ R.T) + t
    return Xc, Xw


def world_to_cam(T, Xw):
    T = T.to(device=Xw.device, dtype=torch.float64)
    R = quat_to_R(T[3:7])
    return ((Xw - T[0:3]) @ R) / T[7]


def project(K, Xc):
    K = K.to(device=Xc.device, dtype=torch.float64)
    u = K[0, 0] * Xc[:, 0] / Xc[:, 2] + K[0, 2]
    v = K[1, 1] * Xc[:, 1] / Xc[:, 2] + K[1, 2]
    return u, v


def _desc_field(Xw, params):
    """Smooth F-channel descriptor attached to the sphere surface; unit norm rows."""
    c = torch.tensor(SPHERE_C, device=Xw.device, dtype=torch.float64)
    dirs = (Xw - c) / SPHERE_R
    freq, phase, amp = (p.to(Xw.device) for p in params)  # [F,M,3], [F,M], [F,M]
    arg = torch.einsum("nc,fmc->nfm", dirs, freq) + phase
    D = (amp * torch.sin(arg)).sum(-1)
    return D / D.norm(dim=-1, keepdim=True)


def _desc_params(F, gen, M=4, freq=60.0):
    fr = torch.randn(F, M, 3, generator=gen, dtype=torch.float64) * freq
    ph = torch.rand(F, M, generator=gen, dtype=torch.float64) * 2 * math.pi
    am = torch.rand(F, M, generator=gen, dtype=torch.float64) + 0.5
    return fr, ph, am


def matching_pairs(B, H, W, seed=0, device="cpu", F=24, desc_noise=0.02, rot_deg=3.0, trans=0.05):
    """B frame pairs.  Returns dict with X11,X21 [B,H,W,3] f32, D11,D21 [B,H,W,F] f32,
    p_true [B,N,2] f64 (analytic sub-pixel match of every image-2 pixel in image 1)."""
    gen = torch.Generator().manual_seed(seed)
    K = tum_intrinsics(H, W, torch.float64)
    T1 = random_poses(B, gen, rot_deg, trans)
    T2 = random_poses(B, gen, rot_deg, trans)
    dp = _desc_params(F, gen)
    out = {k: [] for k in ("X11", "X21", "D11", "D21", "p_true")}
    for b in range(B):
        Xc1, Xw1 = pointmap(K, T1[b], H, W, device)
        _, Xw2 = pointmap(K, T2[b], H, W, device)
        X21 = world_to_cam(T1[b], Xw2)  # image-2 points in camera-1 frame
        u, v = project(K, X21)
        nz = torch.randn(H * W, F, generator=gen, dtype=torch.float64).to(device) * desc_noise
        D11 = _desc_field(Xw1, dp)
        D21 = _desc_field(Xw2, dp) + nz
        D21 = D21 / D21.norm(dim=-1, keepdim=True)
        out["X11"].append(Xc1.float().reshape(H, W, 3))
        out["X21"].append(X21.float().reshape(H, W, 3))
        out["D11"].append(D11.float().reshape(H, W, F))
        out["D21"].append(D21.float().reshape(H, W, F))
        out["p_true"].append(torch.stack([u, v], -1))
    return {k: torch.stack(v).contiguous() for k, v in out.items()}


def img_gradient(img):
    """3x3 gradient used by the reference's matching prep (image.py:5-38): reflect pad,
    kernels [-3,0,3;-10,0,10;-3,0,3]/32 and its transpose, depthwise.  img [b,c,h,w]."""
    import torch.nn.functional as Fn
    c = img.shape[1]
    gx = torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]],
                      device=img.device, dtype=img.dtype) / 32.0
    gy = gx.T.contiguous()
    pad = Fn.pad(img, (1, 1, 1, 1), mode="reflect")
    return (Fn.conv2d(pad, gx.repeat(c, 1, 1, 1), groups=c),
            Fn.conv2d(pad, gy.repeat(c, 1, 1, 1), groups=c))


def prep_for_iter_proj(X11, X21, idx_1_to_2_init=None):
    """Same tensors the reference's matching.py:25-49 hands to iter_proj."""
    import torch.nn.functional as Fn
    b, h, w, _ = X11.shape
    rays = Fn.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    rays_with_grad = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts3d_norm = Fn.normalize(X21.reshape(b, -1, 3), dim=-1).contiguous()
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = torch.stack((idx_1_to_2_init % w, idx_1_to_2_init // w), dim=-1).float().contiguous()
    return rays_with_grad, pts3d_norm, p_init


# --------------------------------------------------------------------------- graphs

def graph_edges(cfg, seed=2):
    """Undirected edge lists of the BASELINE.json configs; returns (K, [(i,j)...])."""
    if cfg == "pair":
        return 2, [(0, 1)]
    if cfg == "window8":  # offsets from tracker.py:31 -> (k-1,k), (k-3,k)
        K = 8
        und = [(k - 1, k) for k in range(1, K)] + [(k - 3, k) for k in range(3, K)]
        return K, und
    if cfg in ("loop64", "calib32") or cfg.startswith("loops:"):
        if cfg.startswith("loops:"):  # "loops:K": K-1 chain edges + K+1 loop closures (E = 4K)
            K = int(cfg.split(":")[1])
            n_loop, min_gap = K + 1, max(2, min(8, K // 8))
        else:
            K, n_loop, min_gap = (64, 65, 8) if cfg == "loop64" else (32, 33, 4)
        und = [(k - 1, k) for k in range(1, K)]
        available = sum(K - g for g in range(min_gap, K))  # pairs (i, j) with j - i >= min_gap
        if n_loop > available:
            raise ValueError(f"{cfg}: {n_loop} loop closures asked, {available} distinct pairs exist")
        gen = torch.Generator().manual_seed(seed)
        seen = set(und)
        while len(und) < K - 1 + n_loop:
            i, j = sorted(torch.randint(0, K, (2,), generator=gen).tolist())
            if j - i >= min_gap and (i, j) not in seen:
                seen.add((i, j))
                und.append((i, j))
        return K, und
    raise ValueError(cfg)


def keyframe_graph(num_kf, H, W, undirected, seed=1, device="cpu", pose_noise=0.01,
                   valid_frac=0.8, id_base=0, id_stride=1, jitter_px=1.0):
    """Keyframe graph with both edge directions (global_opt.py:197-201).

    Returns dict: Twc [K,8] f32 (perturbed), Twc_true, Xs [K,N,3], Cs [K,N,1], K [3,3],
    ii,jj [E] i64 (keyframe ids), idx_ii2jj [E,N] i64, valid_match [E,N,1] bool, Q [E,N,1] f32.
    Edge e=(i,j): for point n of frame j, idx[e,n] is its pixel (linear) in frame i.
    """
    gen = torch.Generator().manual_seed(seed)
    N = H * W
    Kmat = tum_intrinsics(H, W, torch.float64)
    T_true = random_poses(num_kf, gen, rot_deg=4.0, trans=0.08, scale=0.05)
    Xs, Xw = [], []
    for k in range(num_kf):
        xc, xw = pointmap(Kmat, T_true[k], H, W, device)
        Xs.append(xc.float())
        Xw.append(xw)
    Xs = torch.stack(Xs).contiguous()
    Cs = (torch.rand(num_kf, N, 1, generator=gen) * 2.5 + 0.5).to(device)
    directed = [(i, j) for (i, j) in undirected] + [(j, i) for (i, j) in undirected]
    E = len(directed)
    idx = torch.empty(E, N, dtype=torch.int64, device=device)
    valid = torch.empty(E, N, 1, dtype=torch.bool, device=device)
    gen_e = torch.Generator().manual_seed(seed + 2)
    for e, (i, j) in enumerate(directed):
        u, v = project(Kmat, world_to_cam(T_true[i], Xw[j]))
        jit = torch.randn(N, 2, generator=gen_e, dtype=torch.float64).to(device) * jitter_px
        ui = torch.round(u + jit[:, 0])
        vi = torch.round(v + jit[:, 1])
        inside = (ui >= 0) & (ui < W) & (vi >= 0) & (vi < H)
        keep = (torch.rand(N, generator=gen_e) < valid_frac).to(device)
        idx[e] = (ui.clamp(0, W - 1) + W * vi.clamp(0, H - 1)).long()
        valid[e, :, 0] = inside & keep
    Q = (torch.rand(E, N, 1, generator=gen_e) * 2.0 + 1.0).to(device)
    # perturb: T = T_true with tangent-ish noise on t, q, s
    noise = torch.randn(num_kf, 8, generator=gen, dtype=torch.float64) * pose_noise
    T = T_true.clone()
    T[:, 0:3] += noise[:, 0:3]
    T[:, 3:7] += 0.5 * noise[:, 3:7]
    T[:, 3:7] /= T[:, 3:7].norm(dim=-1, keepdim=True)
    T[:, 7] *= torch.exp(noise[:, 7])
    T[0] = T_true[0]
    ids = id_base + id_stride * torch.arange(num_kf)
    ii = torch.tensor([ids[i] for i, _ in directed], dtype=torch.int64, device=device)
    jj = torch.tensor([ids[j] for _, j in directed], dtype=torch.int64, device=device)
    return dict(Twc=T.float().to(device).contiguous(), Twc_true=T_true.float().to(device),
                Xs=Xs, Cs=Cs.contiguous(), K=Kmat.float().to(device).contiguous(), ii=ii, jj=jj,
                idx_ii2jj=idx.contiguous(), valid_match=valid.contiguous(), Q=Q.contiguous(),
                H=H, W=W)


# ----------------------------------------------------------------------- odometry edges

ODOM_PARAMS = dict(sigma_odom_t=0.001, sigma_odom_r=0.001, sigma_ray=0.1,
                   sigma_scale_prior=0.1)  # hard-coded by the caller, global_opt.py:253-257


def _quat_mul(a, b):
    """Hamilton product of [x,y,z,w] quaternions."""
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], dim=-1)


def odometry_edges(g, seed=7, noise_t=2e-4, noise_r=2e-4, skip_prior=(2,), device=None):
    """Odometry inputs of gauss_newton_rays_odom for the keyframe graph `g` (keyframe_graph with
    id_base=0, id_stride=1: the reference uses the ids as Twc rows, gn_kernels.cu:1121-1122):
    edges k -> k+1 with measured relative pose [R_k^T R_{k+1}, R_k^T (t_{k+1} - t_k) / s_{k+1}]
    of the TRUE poses (the prediction of gn_kernels.cu:1146-1155) plus small noise, and scale
    targets s_bar = true scale (0 for the keyframes in `skip_prior`: skipped by the prior).
    Returns dict(odom_ii, odom_jj [K-1] i64, Tij [K-1,8] f32, s_bar [K] f64 on the CPU)."""
    T = g["Twc_true"].double().cpu()
    K = T.shape[0]
    gen = torch.Generator().manual_seed(seed)
    i = torch.arange(0, K - 1)
    j = i + 1
    qi, qj = T[i, 3:7], T[j, 3:7]
    qi_inv = qi * torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=torch.float64)
    q_rel = _quat_mul(qi_inv, qj)
    axis = torch.randn(K - 1, 3, generator=gen, dtype=torch.float64) * noise_r
    dq = torch.cat([0.5 * axis, torch.ones(K - 1, 1, dtype=torch.float64)], dim=-1)
    q_meas = _quat_mul(q_rel, dq)
    q_meas = q_meas / q_meas.norm(dim=-1, keepdim=True)
    Ri = quat_to_R(qi)
    t_rel = torch.einsum("nji,nj->ni", Ri, T[j, 0:3] - T[i, 0:3]) / T[j, 7:8]
    t_meas = t_rel + torch.randn(K - 1, 3, generator=gen, dtype=torch.float64) * noise_t
    Tij = torch.cat([t_meas, q_meas, torch.ones(K - 1, 1, dtype=torch.float64)], dim=-1).float()
    s_bar = T[:, 7].clone()
    for k in skip_prior:
        if k < K:
            s_bar[k] = 0.0
    dev = device if device is not None else g["Twc"].device
    return dict(odom_ii=i.to(dev), odom_jj=j.to(dev), Tij=Tij.contiguous().to(dev), s_bar=s_bar)


# ------------------------------------------------------------------ tracker

# config/base.yaml:18-35 of the reference (the `tracking` section)
TRACK_PARAMS = dict(max_iters=50, C_conf=0.0, Q_conf=1.5, rel_error=1e-3, delta_norm=1e-3,
                    huber=1.345, sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0,
                    sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6)


def tracking_case(H, W, seed=1, device="cpu", pose_noise=0.01, valid_frac=0.8, jitter_px=1.0):
    """One frame tracked against one keyframe (tracker.py:160-300): the inputs of
    FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 at FULL size (no compaction).

    Keyframe = keyframe 0 and frame = keyframe 1 of a two-keyframe graph; the directed edge
    (i = frame, j = keyframe) carries idx_f2k (for keyframe pixel n, its pixel in the frame),
    valid_match and Q.  Returns dict: Xf_all [N,3] (the frame's own pointmap), idx_f2k [N] i64,
    Xf [N,3] = Xf_all[idx_f2k] (tracker.py:395), Xk [N,3], Cf, Ck [N,1], Qk [N,1],
    valid [N,1] bool = valid_match & Q > Q_conf & C > C_conf (tracker.py:199-204),
    T_WCf, T_WCk [1,8] (the frame's pose perturbed), T_WCf_true, K [3,3], H, W."""
    g = keyframe_graph(2, H, W, [(0, 1)], seed=seed, device=device, pose_noise=pose_noise,
                       valid_frac=valid_frac, jitter_px=jitter_px)
    # directed edges are [(0,1), (1,0)]; edge 1 has i = 1 (frame), j = 0 (keyframe):
    # idx[1, n] = pixel of the FRAME matched to keyframe pixel n
    e = 1
    idx = g["idx_ii2jj"][e].contiguous()
    Xf_all = g["Xs"][1].contiguous()
    Xk = g["Xs"][0].contiguous()
    Cf = g["Cs"][1][idx].contiguous()
    Ck = g["Cs"][0].contiguous()
    Qk = g["Q"][e].contiguous()
    tp = TRACK_PARAMS
    valid = (g["valid_match"][e] & (Qk > tp["Q_conf"]) & (Cf > tp["C_conf"])
             & (Ck > tp["C_conf"])).contiguous()
    return dict(Xf_all=Xf_all, idx_f2k=idx, Xf=Xf_all[idx].contiguous(), Xk=Xk, Cf=Cf, Ck=Ck,
                Qk=Qk, valid=valid, T_WCf=g["Twc"][1:2].contiguous(),
                T_WCk=g["Twc"][0:1].contiguous(), T_WCf_true=g["Twc_true"][1:2].contiguous(),
                K=g["K"], H=H, W=W)
