"""ctypes binding of the CPU oracle (oracle/frodo_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package never
imports this module.  All arrays are numpy, C-contiguous.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libfrodo_oracle.so")
_lib = None

_c_i = ctypes.c_int
_c_f = ctypes.c_float
_c_p = ctypes.c_void_p


def build(force=False):
    src = os.path.join(_HERE, "frodo_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libfrodo_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_num_threads.restype = _c_i
        _lib.oracle_edge_index.restype = _c_i
        _lib.oracle_chol_solve.restype = _c_i
        _lib.oracle_gauss_newton.restype = _c_i
    return _lib


def _p(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def _c(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def num_threads():
    return lib().oracle_num_threads()


def set_num_threads(n):
    lib().oracle_set_num_threads(_c_i(n))


def iter_proj(rays, pts, p_init, max_iter, lambda_init, cost_thresh):
    """matching_kernels.cu:128-284.  Returns (p_new f32 [B,N,2], converged bool [B,N])."""
    rays = _c(rays, np.float32)
    pts = _c(pts, np.float32)
    p_init = _c(p_init, np.float32)
    B, H, W, C = rays.shape
    assert C == 9
    N = p_init.shape[1]
    p_new = np.zeros((B, N, 2), np.float32)
    conv = np.zeros((B, N), np.uint8)
    lib().oracle_iter_proj(_p(rays), _p(pts), _p(p_init), _p(p_new), _p(conv), _c_i(B), _c_i(H),
                           _c_i(W), _c_i(N), _c_i(max_iter), _c_f(lambda_init), _c_f(cost_thresh))
    return p_new, conv.astype(bool)


def refine_matches(D11, D21, p1, radius, dilation_max):
    """matching_kernels.cu:25-85.  D11/D21 float16 or float32.  Returns (p1_new i64, scores)."""
    p1 = _c(p1, np.int64)
    B, H, W, F = D11.shape
    N = p1.shape[1]
    p1_new = np.zeros((B, N, 2), np.int64)
    if D11.dtype == np.float16:
        D11 = _c(D11, np.float16)
        D21 = _c(D21, np.float16)
        scores = np.zeros((B, N), np.float16)
        fn = lib().oracle_refine_matches_f16
    elif D11.dtype == np.float32:
        D11 = _c(D11, np.float32)
        D21 = _c(D21, np.float32)
        scores = np.zeros((B, N), np.float32)
        fn = lib().oracle_refine_matches_f32
    else:
        raise TypeError(D11.dtype)
    fn(_p(D11), _p(D21), _p(p1), _p(p1_new), _p(scores), _c_i(B), _c_i(H), _c_i(W), _c_i(N),
       _c_i(F), _c_i(radius), _c_i(dilation_max))
    return p1_new, scores


def edge_index(ii, jj):
    """get_unique_kf_idx + create_inds(pin=0), gn_kernels.cu:203-212."""
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    E = ii.shape[0]
    ie = np.zeros(E, np.int64)
    je = np.zeros(E, np.int64)
    nu = lib().oracle_edge_index(_p(ii), _p(jj), _c_i(E), _p(ie), _p(je))
    return ie, je, nu


MODES = {"rays": 0, "points": 1, "calib": 2}


def edge_blocks(mode, Twc, Xs, Cs, ii_edge, jj_edge, idx, valid, Q, sigma_a, sigma_b, C_thresh,
                Q_thresh, K=None, height=0, width=0, pixel_border=0, z_eps=0.0):
    """Per-edge Hs [4,E,7,7], gs [2,E,7] in fp64 (reference layout, gn_kernels.cu:865-906)."""
    Twc = _c(Twc, np.float32).reshape(-1, 8)
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    nK, N = Xs.shape[0], Xs.shape[1]
    ii_edge = _c(ii_edge, np.int64)
    jj_edge = _c(jj_edge, np.int64)
    E = ii_edge.shape[0]
    idx = _c(idx, np.int64).reshape(E, N)
    valid = _c(np.asarray(valid).reshape(E, N), np.uint8)
    Q = _c(Q, np.float32).reshape(E, N)
    Kc = None if K is None else _c(K, np.float32)
    Hs = np.zeros((4, E, 7, 7), np.float64)
    gs = np.zeros((2, E, 7), np.float64)
    lib().oracle_edge_blocks(_c_i(MODES[mode]), _p(Twc), _p(Xs), _p(Cs), _p(Kc), _p(ii_edge),
                             _p(jj_edge), _p(idx), _p(valid), _p(Q), _c_i(nK), _c_i(N), _c_i(E),
                             _c_i(height), _c_i(width), _c_i(pixel_border), _c_f(z_eps),
                             _c_f(sigma_a), _c_f(sigma_b), _c_f(C_thresh), _c_f(Q_thresh), _p(Hs),
                             _p(gs))
    return Hs, gs


def assemble(Hs, gs, ii_edge, jj_edge, num_kf, num_fix, round_f32=True):
    Hs = _c(Hs, np.float64)
    gs = _c(gs, np.float64)
    ii_edge = _c(ii_edge, np.int64)
    jj_edge = _c(jj_edge, np.int64)
    E = ii_edge.shape[0]
    n = 7 * (num_kf - num_fix)
    H = np.zeros((n, n), np.float64)
    g = np.zeros(n, np.float64)
    lib().oracle_assemble(_p(Hs), _p(gs), _p(ii_edge), _p(jj_edge), _c_i(E), _c_i(num_kf),
                          _c_i(num_fix), _c_i(1 if round_f32 else 0), _p(H), _p(g))
    return H, g


def chol_solve(H, g):
    H = _c(H, np.float64)
    g = _c(g, np.float64)
    n = g.shape[0]
    x = np.zeros(n, np.float64)
    fail = lib().oracle_chol_solve(_p(H), _p(g), _c_i(n), _p(x))
    return x, bool(fail)


def pose_retr(Twc, dx, num_fix):
    Twc = np.array(Twc, dtype=np.float32, copy=True).reshape(-1, 8)
    dx = _c(dx, np.float32)
    lib().oracle_pose_retr(_p(Twc), _p(dx), _c_i(Twc.shape[0]), _c_i(num_fix))
    return Twc


def gauss_newton(mode, Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_a, sigma_b, C_thresh, Q_thresh,
                 num_fix, max_iter, delta_thresh, K=None, height=0, width=0, pixel_border=0,
                 z_eps=0.0, return_system=False):
    """gauss_newton_{rays,points,calib}_cuda (gn_kernels.cu:909-995, 2206-2293, 2610-2702).

    Returns dict(Twc=updated copy, dx=[K-num_fix,7], iters=int, fails=int[, H, g of iter 0]).
    """
    Twc = np.array(Twc, dtype=np.float32, copy=True).reshape(-1, 8)
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    nK, N = Xs.shape[0], Xs.shape[1]
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    E = ii.shape[0]
    idx = _c(idx, np.int64).reshape(E, N)
    valid = _c(np.asarray(valid).reshape(E, N), np.uint8)
    Q = _c(Q, np.float32).reshape(E, N)
    Kc = None if K is None else _c(K, np.float32)
    n = 7 * (nK - num_fix)
    dx = np.zeros((nK - num_fix, 7), np.float32)
    H = np.zeros((n, n), np.float64) if return_system else None
    g = np.zeros(n, np.float64) if return_system else None
    fails = ctypes.c_int(0)
    iters = lib().oracle_gauss_newton(
        _c_i(MODES[mode]), _p(Twc), _p(Xs), _p(Cs), _p(Kc), _p(ii), _p(jj), _p(idx), _p(valid),
        _p(Q), _c_i(nK), _c_i(N), _c_i(E), _c_i(height), _c_i(width), _c_i(pixel_border),
        _c_f(z_eps), _c_f(sigma_a), _c_f(sigma_b), _c_f(C_thresh), _c_f(Q_thresh), _c_i(num_fix),
        _c_i(max_iter), _c_f(delta_thresh), _p(dx), _p(H), _p(g), ctypes.byref(fails))
    out = dict(Twc=Twc, dx=dx, iters=iters, fails=fails.value)
    if return_system:
        out["H"] = H
        out["g"] = g
    return out


def odom_blocks(Twc, Tij, odom_ii, odom_jj, sigma_odom_t, sigma_odom_r):
    """odom_constraint_kernel_left_perturb_log (gn_kernels.cu:1106-1286): Hs_odom [4,Eo,7,7],
    gs_odom [2,Eo,7] in fp32.  odom_ii / odom_jj index Twc rows directly (:1121-1122)."""
    Twc = _c(Twc, np.float32).reshape(-1, 8)
    Tij = _c(Tij, np.float32)
    Tij = Tij.reshape(Tij.shape[0], -1)
    odom_ii = _c(odom_ii, np.int64)
    odom_jj = _c(odom_jj, np.int64)
    Eo = odom_ii.shape[0]
    Hs = np.zeros((4, Eo, 7, 7), np.float32)
    gs = np.zeros((2, Eo, 7), np.float32)
    if Eo:
        lib().oracle_odom_blocks(_p(Twc), _p(Tij), _c_i(Tij.shape[1]), _p(odom_ii), _p(odom_jj),
                                 _c_i(Eo), _c_f(sigma_odom_t), _c_f(sigma_odom_r), _p(Hs), _p(gs))
    return Hs, gs


def add_blocks(H, g, Hs, gs, i_opt, j_opt):
    """SparseBlock::update_lhs / update_rhs (gn_kernels.cu:72-122) for blocks [4,E,7,7] /
    [2,E,7] with system block indices i_opt, j_opt (negative = fixed, skipped)."""
    rows = [i_opt, i_opt, j_opt, j_opt]
    cols = [i_opt, j_opt, i_opt, j_opt]
    for b in range(4):
        for e in range(Hs.shape[1]):
            i, j = int(rows[b][e]), int(cols[b][e])
            if i >= 0 and j >= 0:
                H[7 * i:7 * i + 7, 7 * j:7 * j + 7] += Hs[b, e].astype(np.float64)
    for b, ind in enumerate((i_opt, j_opt)):
        for e in range(gs.shape[1]):
            i = int(ind[e])
            if i >= 0:
                g[7 * i:7 * i + 7] += gs[b, e].astype(np.float64)


def scale_prior(H, g, Twc, s_bar, num_fix, sigma_scale_prior):
    """apply_scale_prior_cpu (gn_kernels.cu:1716-1790): residual log(s_i / sbar_i), J = 1."""
    if sigma_scale_prior <= 1e-9:
        return
    info = 1.0 / (float(sigma_scale_prior) * float(sigma_scale_prior))
    Twc64 = np.asarray(Twc, np.float32).reshape(-1, 8).astype(np.float64)
    s_bar = np.asarray(s_bar, np.float64)
    for i in range(num_fix, Twc64.shape[0]):
        if s_bar[i] <= 1e-6 or Twc64[i, 6] <= 1e-6:  # sic: A.M - 1 = 6 indexes q_w, not s (:1736,1762)
            continue
        k = 7 * (i - num_fix) + 6
        H[k, k] += info
        g[k] += info * np.log(Twc64[i, 6] / s_bar[i])


def gauss_newton_rays_odom(Twc, Xs, Cs, ii, jj, idx, valid, Q, odom_ii, odom_jj, Tij, s_bar,
                           sigma_odom_t, sigma_odom_r, sigma_ray, sigma_dist, sigma_scale_prior,
                           C_thresh, Q_thresh, num_fix, max_iter, delta_thresh):
    """gauss_newton_rays_odom_cuda (gn_kernels.cu:1792-1934), composed from the pieces above.
    Returns dict(Twc, dx, iters, fails)."""
    Twc = np.array(Twc, dtype=np.float32, copy=True).reshape(-1, 8)
    nK = np.asarray(Xs).shape[0]
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    odom_ii = _c(odom_ii, np.int64)
    odom_jj = _c(odom_jj, np.int64)
    ie, je, _ = edge_index(ii, jj)
    uniq = np.unique(np.concatenate([ii, jj]))
    oi_opt = np.searchsorted(uniq, odom_ii) - num_fix  # create_inds(unique, num_fix, odom_ii, ..)
    oj_opt = np.searchsorted(uniq, odom_jj) - num_fix
    n = 7 * (nK - num_fix)
    dx = np.zeros((nK - num_fix, 7), np.float32)
    iters = fails = 0
    for _ in range(max_iter):
        Hs, gs = edge_blocks("rays", Twc, Xs, Cs, ie, je, idx, valid, Q, sigma_ray, sigma_dist,
                             C_thresh, Q_thresh)
        H, g = assemble(Hs, gs, ie, je, nK, num_fix, round_f32=True)
        if odom_ii.shape[0] > 0:
            Ho, go = odom_blocks(Twc, Tij, odom_ii, odom_jj, sigma_odom_t, sigma_odom_r)
            add_blocks(H, g, Ho, go, oi_opt, oj_opt)
        scale_prior(H, g, Twc, s_bar, num_fix, sigma_scale_prior)
        x, fail = chol_solve(H, g)
        fails += int(fail)
        dx = (-x).astype(np.float32).reshape(nK - num_fix, 7)
        Twc = pose_retr(Twc, dx, num_fix)
        iters += 1
        if np.float32(np.sqrt(np.sum(dx.astype(np.float64) ** 2))) < delta_thresh:
            break
    return dict(Twc=Twc, dx=dx, iters=iters, fails=fails)


# ---------------------------------------------------------------------------
# The matching pipeline around the two kernels: restatement of
# mast3r_slam/matching.py:25-110 and mast3r_slam/image.py:5-38 in numpy fp32 with a FIXED
# operation order (every product and sum rounded separately, no FMA), which the CUDA kernels of
# csrc/match_glue.cu follow bit for bit.  torch's own kernels use other orders, so against the
# reference's outputs (tests/golden/golden_match.npz) the prologue agrees to 1-2 ulp, not bitwise.
# ---------------------------------------------------------------------------
_F = np.float32


def normalize3(x):
    """F.normalize(x, dim=-1) (matching.py:30,42): x / max(||x||_2, 1e-12)."""
    x = np.asarray(x, _F)
    n2 = ((x[..., 0] * x[..., 0] + x[..., 1] * x[..., 1]).astype(_F)
          + x[..., 2] * x[..., 2]).astype(_F)
    d = np.maximum(np.sqrt(n2).astype(_F), _F(1e-12))
    return (x / d[..., None]).astype(_F)


def img_gradient(img):
    """image.py:5-38 on a channels-last image [b,h,w,c]: Scharr kernels / 32, reflect padding.
    Taps are accumulated in row-major order, zero taps skipped."""
    img = np.asarray(img, _F)
    h, w = img.shape[1], img.shape[2]
    pad = np.pad(img, ((0, 0), (1, 1), (1, 1), (0, 0)), mode="reflect")
    w3, w10 = _F(3.0 / 32.0), _F(10.0 / 32.0)

    def tap(dy, dx):
        return pad[:, dy:dy + h, dx:dx + w, :]

    gx = np.zeros_like(img)
    for dy, dx, wt in ((0, 0, -w3), (0, 2, w3), (1, 0, -w10), (1, 2, w10), (2, 0, -w3), (2, 2, w3)):
        gx = (gx + (tap(dy, dx) * wt).astype(_F)).astype(_F)
    gy = np.zeros_like(img)
    for dy, dx, wt in ((0, 0, -w3), (0, 1, -w10), (0, 2, -w3), (2, 0, w3), (2, 1, w10), (2, 2, w3)):
        gy = (gy + (tap(dy, dx) * wt).astype(_F)).astype(_F)
    return gx, gy


def prep_for_iter_proj(X11, X21, idx_1_to_2_init=None):
    """matching.py:25-49.  Returns rays_with_grad_img [b,h,w,9], pts3d_norm [b,hw,3],
    p_init [b,hw,2] (x, y) float32."""
    X11 = np.asarray(X11, _F)
    b, h, w, _ = X11.shape
    rays = normalize3(X11)
    gx, gy = img_gradient(rays)
    rays_with_grad = np.ascontiguousarray(np.concatenate([rays, gx, gy], axis=-1))
    pts = normalize3(np.asarray(X21, _F).reshape(b, h * w, 3))
    if idx_1_to_2_init is None:
        idx_1_to_2_init = np.tile(np.arange(h * w, dtype=np.int64)[None], (b, 1))
    idx0 = np.asarray(idx_1_to_2_init, np.int64)
    p_init = np.stack([idx0 % w, idx0 // w], axis=-1).astype(_F)
    return rays_with_grad, pts, p_init


def match_occlusion(X11, X21, p_iter, conv, dist_thresh):
    """matching.py:66-75: p1 = trunc(p), valid = converged & ||X11[p1] - X21|| < dist_thresh."""
    X11 = np.asarray(X11, _F)
    b, h, w, _ = X11.shape
    p1 = np.trunc(np.asarray(p_iter, _F)).astype(np.int64)
    bi = np.arange(b)[:, None]
    d = (X11[bi, p1[..., 1], p1[..., 0], :] - np.asarray(X21, _F).reshape(b, h * w, 3)).astype(_F)
    n2 = ((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(_F)
          + d[..., 2] * d[..., 2]).astype(_F)
    dist = np.sqrt(n2).astype(_F)
    return p1, np.asarray(conv, bool) & (dist < _F(dist_thresh))


def match_finish(p1, scores, valid, w, ft_matching_threshold):
    """matching.py:87-92: valid &= score > threshold (compared in the scores' dtype),
    idx = x + w * y.  Returns (idx [b,hw] int64, valid [b,hw,1] bool)."""
    scores = np.asarray(scores)
    thr = scores.dtype.type(ft_matching_threshold)
    v = np.asarray(valid, bool) & (scores > thr)
    p1 = np.asarray(p1, np.int64)
    return p1[..., 0] + w * p1[..., 1], v[..., None]


def match(X11, X21, D11, D21, idx_1_to_2_init=None, max_iter=10, lambda_init=1e-8,
          convergence_thresh=1e-6, dist_thresh=0.1, radius=3, dilation_max=5,
          ft_matching_threshold=0.8):
    """matching.py:8-10, 52-95 (match -> match_iterative_proj).  D11/D21 any float dtype; they
    are rounded to fp16 as the reference does (matching.py:79-80)."""
    X11 = np.asarray(X11, _F)
    b, h, w, _ = X11.shape
    rays, pts, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p, conv = iter_proj(rays, pts, p_init, max_iter, lambda_init, convergence_thresh)
    p1, valid = match_occlusion(X11, X21, p, conv, dist_thresh)
    D11h = np.asarray(D11).astype(np.float16)
    D21h = np.asarray(D21).astype(np.float16).reshape(b, h * w, -1)
    p1, scores = refine_matches(D11h, D21h, p1, radius, dilation_max)
    return match_finish(p1, scores, valid, w, ft_matching_threshold)


# ----------------------------------------------------------------------------------------
# The tracker's frame-to-keyframe pose solve: tracker.py:397-517 with
# geometry.py:17-34 (point_to_ray_dist), :45-52 (act_Sim3), :63-104 (project_calib) and
# nonlinear_optimizer.py:4-33 (check_convergence, huber).  The reference runs it as ~15 PyTorch
# kernels per iteration on fp32 tensors with lietorch.Sim3 group operations (lietorch is an
# unpinned git dependency, absent from /root/reference: its Sim3 act / inv / mul / retr are
# restated here from the published formulas, the same ones gn_kernels.cu:219-314,341-455
# carries; Exp goes through the oracle's pose_retr, pinned to the reference kernel).
# numpy fp32, vectorised over the points.  Pinned against the reference's own tracker.py run
# on a stub lietorch: tests/golden/make_golden_track.py -> golden_track.npz.
# ----------------------------------------------------------------------------------------

def _f32(a):
    return np.asarray(a, dtype=np.float32)


def sim3_act(T, X):
    """lietorch Sim3.act: s R x + t.  T [8] = [t, q_xyzw, s], X [N,3]."""
    T = _f32(T).reshape(8)
    X = _f32(X)
    q = T[3:6]
    uv = np.float32(2.0) * np.cross(q[None, :], X).astype(np.float32)
    Y = X + T[6] * uv + np.cross(q[None, :], uv).astype(np.float32)
    return (T[7] * Y + T[None, 0:3]).astype(np.float32)


def _quat_mul(a, b):
    return np.array([a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1],
                     a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0],
                     a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3],
                     a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]], dtype=np.float32)


def sim3_inv(T):
    """lietorch Sim3.inv: [ -s^-1 R^T t, q^-1, s^-1 ]."""
    T = _f32(T).reshape(8)
    qi = np.array([-T[3], -T[4], -T[5], T[6]], np.float32)
    si = np.float32(1.0) / T[7]
    Ti = np.zeros(8, np.float32)
    Ti[3:7] = qi
    Ti[7] = 1.0
    rt = sim3_act(Ti, T[None, 0:3])[0]
    Ti[0:3] = -si * rt
    Ti[7] = si
    return Ti


def sim3_mul(A, B):
    """lietorch Sim3 product A * B: [ s_a R_a t_b + t_a, q_a q_b, s_a s_b ]."""
    A = _f32(A).reshape(8)
    B = _f32(B).reshape(8)
    out = np.zeros(8, np.float32)
    out[0:3] = sim3_act(A, B[None, 0:3])[0]
    out[3:7] = _quat_mul(A[3:7], B[3:7])
    out[7] = A[7] * B[7]
    return out


def sim3_retr(T, tau):
    """lietorch Sim3.retr: Exp(tau) * T (expSim3 / retrSim3, gn_kernels.cu:365-455)."""
    return pose_retr(_f32(T).reshape(1, 8), _f32(tau).reshape(1, 7), 0)[0]


def _huber_w(r, k):
    a = np.abs(r)
    with np.errstate(divide="ignore"):
        return np.where(a < k, np.float32(1.0), np.float32(k) / a).astype(np.float32)


def _track_loop(T_CkCf, T_WCk, step_fn, huber_k, max_iters, rel_error, delta_norm):
    """The iteration of tracker.py:433-460 / :477-512 with FrameTracker.solve (:397-412)."""
    old_cost = float("inf")
    costs = []
    iters = 0
    status = 0
    for step in range(max_iters):
        sqrt_info, r, J = step_fn(T_CkCf)
        whitened = sqrt_info * r
        rsi = (sqrt_info * np.sqrt(_huber_w(whitened, huber_k))).astype(np.float32)
        mdim = J.shape[-1]
        A = (rsi[..., None] * J).reshape(-1, mdim).astype(np.float32)
        b = (rsi * r).reshape(-1, 1).astype(np.float32)
        H = (A.T @ A).astype(np.float32)
        g = (-A.T @ b).astype(np.float32)
        cost = 0.5 * float((b.T @ b)[0, 0])
        try:
            L = np.linalg.cholesky(H.astype(np.float64))
        except np.linalg.LinAlgError:
            status = 1  # torch.linalg.cholesky raises; track() reports "Cholesky failed"
            break
        tau = np.linalg.solve(L.T, np.linalg.solve(L, g.astype(np.float64))).astype(np.float32)
        T_CkCf = sim3_retr(T_CkCf, tau.reshape(7))
        costs.append(cost)
        iters = step + 1
        # check_convergence, nonlinear_optimizer.py:4-25 (old_cost = inf on step 0 -> nan -> False)
        with np.errstate(invalid="ignore"):
            rel_dec = abs(float(np.float64(old_cost - cost) / np.float64(old_cost))) \
                if np.isfinite(old_cost) else float("nan")
        dn = float(np.linalg.norm(tau))
        if rel_dec < rel_error or dn < delta_norm:
            break
        old_cost = cost
    T_WCf = sim3_mul(T_WCk, T_CkCf)
    return dict(T_WCf=T_WCf, T_CkCf=T_CkCf, iters=iters, costs=np.array(costs, np.float64),
                status=status)


def _act_jac(T, Xf):
    """act_Sim3 with jacobian, geometry.py:45-52: [I, -[p]x, p]."""
    P = sim3_act(T, Xf)
    n = P.shape[0]
    J = np.zeros((n, 3, 7), np.float32)
    J[:, 0, 0] = J[:, 1, 1] = J[:, 2, 2] = 1.0
    x, y, z = P[:, 0], P[:, 1], P[:, 2]
    J[:, 0, 4] = z; J[:, 0, 5] = -y
    J[:, 1, 3] = -z; J[:, 1, 5] = x
    J[:, 2, 3] = y; J[:, 2, 4] = -x
    J[:, :, 6] = P
    return P, J


def _ray_dist(X, jacobian):
    """point_to_ray_dist, geometry.py:17-34."""
    d = np.linalg.norm(X, axis=-1, keepdims=True).astype(np.float32)
    d_inv = (np.float32(1.0) / d).astype(np.float32)
    r = d_inv * X
    rd = np.concatenate((r, d), axis=-1).astype(np.float32)
    if not jacobian:
        return rd
    d_inv_2 = d_inv ** 2
    I = np.eye(3, dtype=np.float32)[None]
    dr = d_inv[..., None] * (I - d_inv_2[..., None] * (X[:, :, None] * X[:, None, :]))
    drd = np.concatenate((dr, r[:, None, :]), axis=-2).astype(np.float32)
    return rd, drd


def track_pose_rays(Xf, Xk, T_WCf, T_WCk, Qk, valid, sigma_ray=0.003, sigma_dist=10.0,
                    huber=1.345, max_iters=50, rel_error=1e-3, delta_norm=1e-3, idx_f2k=None):
    """FrameTracker.opt_pose_ray_dist_sim3, tracker.py:414-465, on FULL-size arrays: the caller's
    compaction by valid_opt (tracker.py:293-295) happens here.  Xf [N,3] is the frame's pointmap
    (already gathered by idx_f2k, or gathered here when idx_f2k is given, tracker.py:395),
    Xk [N,3] the keyframe's, Qk [N,1], valid [N,1] bool, poses [8]."""
    Xf = _f32(Xf).reshape(-1, 3)
    if idx_f2k is not None:
        Xf = Xf[np.asarray(idx_f2k).reshape(-1)]
    m = np.asarray(valid).reshape(-1).astype(bool)
    Xf = Xf[m]
    Xk = _f32(Xk).reshape(-1, 3)[m]
    Qk = _f32(Qk).reshape(-1, 1)[m]
    sq = np.sqrt(Qk).astype(np.float32)
    si_ray = (np.float32(1.0 / sigma_ray) * sq).astype(np.float32)
    si_dist = (np.float32(1.0 / sigma_dist) * sq).astype(np.float32)
    sqrt_info = np.concatenate((np.repeat(si_ray, 3, axis=1), si_dist), axis=1)
    T_CkCf = sim3_mul(sim3_inv(T_WCk), T_WCf)
    rd_k = _ray_dist(Xk, False)

    def step_fn(T):
        P, dP = _act_jac(T, Xf)
        rd_f, drd = _ray_dist(P, True)
        r = rd_k - rd_f
        J = -(drd @ dP)
        return sqrt_info, r, J.astype(np.float32)

    return _track_loop(T_CkCf, T_WCk, step_fn, huber, max_iters, rel_error, delta_norm)


def track_pose_calib(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                     sigma_pixel=1.0, sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6,
                     huber=1.345, max_iters=50, rel_error=1e-3, delta_norm=1e-3, idx_f2k=None):
    """FrameTracker.opt_pose_calib_sim3, tracker.py:467-517 (+ project_calib, geometry.py:63-104).
    All arrays full size; `valid` multiplies the information (tracker.py:471-472)."""
    Xf = _f32(Xf).reshape(-1, 3)
    if idx_f2k is not None:
        Xf = Xf[np.asarray(idx_f2k).reshape(-1)]
    v = np.asarray(valid).reshape(-1, 1).astype(np.float32)
    Qk = _f32(Qk).reshape(-1, 1)
    meas_k = _f32(meas_k).reshape(-1, 3)
    vmk = np.asarray(valid_meas_k).reshape(-1, 1).astype(bool)
    K = _f32(K).reshape(3, 3)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    sq = np.sqrt(Qk).astype(np.float32)
    si_p = (np.float32(1.0 / sigma_pixel) * v * sq).astype(np.float32)
    si_d = (np.float32(1.0 / sigma_depth) * v * sq).astype(np.float32)
    sqrt_info = np.concatenate((np.repeat(si_p, 2, axis=1), si_d), axis=1)
    T_CkCf = sim3_mul(sim3_inv(T_WCk), T_WCf)
    h, w = img_size

    def step_fn(T):
        P, dP = _act_jac(T, Xf)
        x, y, z = P[:, 0:1], P[:, 1:2], P[:, 2:3]
        with np.errstate(divide="ignore", invalid="ignore"):
            p = (P @ K.T).astype(np.float32)
            p = (p / p[:, 2:3])[:, :2]
            u, vv = p[:, 0:1], p[:, 1:2]
            valid_u = (u > pixel_border) & (u < w - 1 - pixel_border)
            valid_v = (vv > pixel_border) & (vv < h - 1 - pixel_border)
            valid_z = z > depth_eps
            vproj = valid_u & valid_v & valid_z
            logz = np.log(z).astype(np.float32)
            logz[~valid_z] = 0.0
            pz = np.concatenate((p, logz), axis=-1)
            z_inv = (np.float32(1.0) / z[:, 0]).astype(np.float32)
            d = np.zeros((P.shape[0], 3, 3), np.float32)
            d[:, 0, 0] = fx
            d[:, 1, 1] = fy
            d[:, 0, 2] = -fx * x[:, 0] * z_inv
            d[:, 1, 2] = -fy * y[:, 0] * z_inv
            d *= z_inv[:, None, None]
            d[:, 2, 2] = z_inv
            valid2 = vproj & vmk
            si2 = np.where(valid2, sqrt_info, np.float32(0.0)).astype(np.float32)
            r = (meas_k - pz).astype(np.float32)
            J = -(d @ dP)
            # rows with zero information contribute 0 in the reference unless they hold inf / nan;
            # the synthetic inputs keep them finite, and the masks below make that explicit
            bad = ~np.isfinite(r).all(axis=1) | ~np.isfinite(J).all(axis=(1, 2))
            r[bad] = 0.0
            J[bad] = 0.0
        return si2, r, J.astype(np.float32)

    return _track_loop(T_CkCf, T_WCk, step_fn, huber, max_iters, rel_error, delta_norm)


def track_meas_calib(Xf, Xk, K, img_size, depth_eps=1e-6):
    """The calibrated branch of FrameTracker.get_points_poses, tracker.py:383-393 (+
    constrain_points_to_ray / backproject / get_pixel_coords, geometry.py:37-42,107-124): both
    pointmaps are re-projected onto their own pixel rays at their depth; the measurement of the
    keyframe is (u, v, log z) with invalid-depth rows zeroed."""
    h, w = img_size
    K = _f32(K).reshape(3, 3)
    vv, uu = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32),
                         indexing="ij")
    uv = np.stack((uu, vv), axis=-1).reshape(-1, 2)

    def constrain(X):
        X = _f32(X).reshape(-1, 3)
        z = X[:, 2:3]
        t1 = (uv[:, 0] - K[0, 2]) / K[0, 0]
        t2 = (uv[:, 1] - K[1, 2]) / K[1, 1]
        d = np.stack((t1, t2, np.ones_like(t1)), axis=-1).astype(np.float32)
        return (z * d).astype(np.float32)

    Xf2 = constrain(Xf)
    Xk2 = constrain(Xk)
    with np.errstate(divide="ignore", invalid="ignore"):
        meas = np.concatenate((uv, np.log(Xk2[:, 2:3])), axis=-1).astype(np.float32)
    vmk = Xk2[:, 2:3] > depth_eps
    meas[~np.repeat(vmk, 3, axis=1)] = 0.0
    return Xf2, Xk2, meas, vmk


def match_quality(Q_src, Q_dst, idx, valid_match, Q_conf, C_src=None, C_dst=None, C_conf=0.0):
    """tracker.py:194-206 / global_opt.py:98-112 restated in numpy fp32 (batched [B,N]):
    Q = sqrt(Q_src[b, idx] * Q_dst); valid_q = valid_match & (Q > Q_conf); valid_opt = valid_q &
    (C_src[b, idx] > C_conf) & (C_dst > C_conf); counts = (sum valid_opt, sum valid_q) per b."""
    idx = np.asarray(idx, np.int64)
    B = idx.shape[0]
    N = idx.reshape(B, -1).shape[1]
    idx = idx.reshape(B, N)
    qs = _f32(Q_src).reshape(B, -1)
    qd = _f32(Q_dst).reshape(B, N)
    vm = np.asarray(valid_match).reshape(B, N).astype(bool)
    rows = np.arange(B)[:, None]
    Q = np.sqrt((qs[rows, idx] * qd).astype(np.float32)).astype(np.float32)
    vq = vm & (Q > np.float32(Q_conf))
    vo = vq.copy()
    if C_src is not None:
        vo &= _f32(C_src).reshape(B, -1)[rows, idx] > np.float32(C_conf)
    if C_dst is not None:
        vo &= _f32(C_dst).reshape(B, N) > np.float32(C_conf)
    counts = np.stack((vo.sum(axis=1), vq.sum(axis=1)), axis=1).astype(np.int32)
    return Q, vo, vq, counts

