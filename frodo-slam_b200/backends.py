"""The reference's ``mast3r_slam_backends`` operator surface on top of libfrodo_b200.

Same names, positional signatures, return lists and in-place ``Twc`` semantics as
the pybind module of the reference (mast3r_slam/backend/src/gn.cpp:152-160), so
matching.py:60,79, global_opt.py:260,318 and local_mapping.py:172,248 can call it
unchanged.  Differences, all stricter: dtypes/devices are validated, kernels run
on torch's *current* stream with the tensors' device made current, and the
Gauss-Newton drivers never synchronise with the host.

Host code here only validates, allocates outputs with torch and forwards raw
device pointers to the C ABI (include/frodo_b200.h).  No CPU fallback exists.
"""
import ctypes

import torch

from . import _lib
from ._lib import GnParams, GN_RAYS, GN_POINTS, GN_CALIB

__all__ = ["iter_proj", "refine_matches", "gauss_newton_rays", "gauss_newton_points",
           "gauss_newton_calib", "gauss_newton_rays_odom", "gn_system"]


def _check_contiguous(**tensors):
    # gn.h:5  TORCH_CHECK(x.is_contiguous(), #x " must be contiguous") -> RuntimeError
    for name, t in tensors.items():
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")


def _check(t, name, dtype, device=None):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (frodo-slam_b200 has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    if device is not None and t.device != device:
        raise RuntimeError(f"{name} is on {t.device}, expected {device}")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    """gn.cpp:120-135 / matching_kernels.cu:288-325.  Returns [p_new f32 [B,N,2], converged bool [B,N]]."""
    _check_contiguous(rays_img_with_grad=rays_img_with_grad, pts_3d_norm=pts_3d_norm,
                      p_init=p_init)
    dev = p_init.device
    _check(rays_img_with_grad, "rays_img_with_grad", torch.float32)
    _check(pts_3d_norm, "pts_3d_norm", torch.float32, dev)
    _check(p_init, "p_init", torch.float32, dev)
    if rays_img_with_grad.dim() != 4 or rays_img_with_grad.shape[3] != 9:
        raise RuntimeError("rays_img_with_grad must be [B,H,W,9]")
    B, H, W, _ = rays_img_with_grad.shape
    N = p_init.shape[1]
    if p_init.shape != (B, N, 2) or pts_3d_norm.shape != (B, N, 3):
        raise RuntimeError("p_init must be [B,N,2] and pts_3d_norm [B,N,3]")
    lib = _lib.load()
    with torch.cuda.device(dev):
        p_new = torch.empty((B, N, 2), dtype=torch.float32, device=dev)
        converged = torch.empty((B, N), dtype=torch.bool, device=dev)
        _lib.check(lib.frodo_iter_proj(_ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init),
                                       _ptr(p_new), _ptr(converged), B, H, W, N, int(max_iter),
                                       float(lambda_init), float(cost_thresh), _stream()))
    return [p_new, converged]


def refine_matches(D11, D21, p1, radius, dilation_max):
    """gn.cpp:137-150 / matching_kernels.cu:88-125.  Returns [p1_new i64 [B,N,2], max_scores [B,N]]."""
    _check_contiguous(D11=D11, D21=D21, p1=p1)
    dev = p1.device
    if D11.dtype not in (torch.float16, torch.float32, torch.float64):
        raise RuntimeError(f"refine_matches supports float16/32/64 descriptors, got {D11.dtype}")
    _check(D11, "D11", D11.dtype, dev)
    _check(D21, "D21", D11.dtype, dev)
    _check(p1, "p1", torch.int64)
    if D11.dim() != 4 or D21.dim() != 3 or p1.dim() != 3:
        raise RuntimeError("expected D11 [B,H,W,F], D21 [B,N,F], p1 [B,N,2]")
    B, H, W, F = D11.shape
    N = p1.shape[1]
    if D21.shape != (B, N, F) or p1.shape != (B, N, 2):
        raise RuntimeError("expected D11 [B,H,W,F], D21 [B,N,F], p1 [B,N,2]")
    lib = _lib.load()
    fn = {torch.float16: lib.frodo_refine_matches_f16, torch.float32: lib.frodo_refine_matches_f32,
          torch.float64: lib.frodo_refine_matches_f64}[D11.dtype]
    with torch.cuda.device(dev):
        p1_new = torch.empty((B, N, 2), dtype=torch.int64, device=dev)
        scores = torch.empty((B, N), dtype=D11.dtype, device=dev)
        _lib.check(fn(_ptr(D11), _ptr(D21), _ptr(p1), _ptr(p1_new), _ptr(scores), B, H, W, N, F,
                      int(radius), int(dilation_max), _stream()))
    return [p1_new, scores]


def _gn_validate(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, K=None):
    named = dict(Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx_ii2jj=idx_ii2jj,
                 valid_match=valid_match, Q=Q)
    if K is not None:
        named["K"] = K
    _check_contiguous(**named)
    dev = Twc.device
    _check(Twc, "Twc", torch.float32)
    _check(Xs, "Xs", torch.float32, dev)
    _check(Cs, "Cs", torch.float32, dev)
    _check(ii, "ii", torch.int64, dev)
    _check(jj, "jj", torch.int64, dev)
    _check(idx_ii2jj, "idx_ii2jj", torch.int64, dev)
    _check(valid_match, "valid_match", torch.bool, dev)
    _check(Q, "Q", torch.float32, dev)
    if K is not None:
        _check(K, "K", torch.float32, dev)
        if K.numel() != 9:
            raise RuntimeError("K must be [3,3]")
    num_kf, N = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    if Xs.dim() != 3 or Xs.shape[2] != 3:
        raise RuntimeError("Xs must be [K,N,3]")
    if Twc.numel() != num_kf * 8:
        raise RuntimeError("Twc must hold K x 8 values [t, q_xyzw, s]")
    if Cs.numel() != num_kf * N:
        raise RuntimeError("Cs must be [K,N,1]")
    if jj.shape[0] != E or idx_ii2jj.numel() != E * N or valid_match.numel() != E * N \
            or Q.numel() != E * N:
        raise RuntimeError("ii, jj [E]; idx_ii2jj, valid_match, Q [E,N(,1)] expected")
    return dev, num_kf, N, E


def _gauss_newton(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, num_fix, max_iter,
                  delta_thresh):
    dev, num_kf, N, E = _gn_validate(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, K)
    lib = _lib.load()
    with torch.cuda.device(dev):
        dx = torch.zeros((max(num_kf - num_fix, 0), 7), dtype=torch.float32, device=dev)
        if max_iter <= 0 or num_kf - num_fix <= 0:
            return [dx]
        nbytes = lib.frodo_gn_workspace_bytes(num_kf, E, N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.frodo_gauss_newton(
            ctypes.byref(params), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(K), _ptr(ii), _ptr(jj),
            _ptr(idx_ii2jj), _ptr(valid_match), _ptr(Q), num_kf, N, E, int(num_fix),
            int(max_iter), float(delta_thresh), _ptr(dx), None, _ptr(ws), nbytes, _stream()))
    return [dx]


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist,
                      C_thresh, Q_thresh, num_fix, max_iter, delta_thresh):
    """gn.cpp:63-88 / gn_kernels.cu:909-995.  Updates Twc in place, returns [dx]."""
    p = GnParams(GN_RAYS, float(sigma_ray), float(sigma_dist), float(C_thresh), float(Q_thresh),
                 0, 0, 0, 0.0)
    return _gauss_newton(p, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, num_fix,
                         max_iter, delta_thresh)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh,
                        Q_thresh, max_iter, delta_thresh):
    """gn.cpp:3-26 / gn_kernels.cu:2206-2293 (num_fix = 1 hard-coded, :2223)."""
    p = GnParams(GN_POINTS, float(sigma_point), 0.0, float(C_thresh), float(Q_thresh), 0, 0, 0, 0.0)
    return _gauss_newton(p, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, 1, max_iter,
                         delta_thresh)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width,
                       pixel_border, z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter,
                       delta_thresh):
    """gn.cpp:90-118 / gn_kernels.cu:2610-2702 (num_fix = 1 hard-coded, :2630)."""
    p = GnParams(GN_CALIB, float(sigma_pixel), float(sigma_depth), float(C_thresh),
                 float(Q_thresh), int(height), int(width), int(pixel_border), float(z_eps))
    return _gauss_newton(p, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, 1, max_iter,
                         delta_thresh)


def gauss_newton_rays_odom(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, odom_ii, odom_jj, Tij,
                           s_bar, sigma_odom_t, sigma_odom_r, sigma_ray, sigma_dist,
                           sigma_scale_prior, C_thresh, Q_thresh, num_fix, max_iter, delta_thresh):
    """gn.cpp:29-62 / gn_kernels.cu:1792-1934: gauss_newton_rays + odometry constraints
    (odom_ii -> odom_jj with measured relative poses Tij [Eo, 7|8]) + scale prior towards s_bar
    [K] (host or device tensor, as the reference's caller passes it, global_opt.py:258).
    Updates Twc in place, returns [dx]."""
    dev, num_kf, N, E = _gn_validate(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, None)
    _check_contiguous(odom_ii=odom_ii, odom_jj=odom_jj, Tij=Tij)
    _check(odom_ii, "odom_ii", torch.int64, dev)
    _check(odom_jj, "odom_jj", torch.int64, dev)
    Eo = odom_ii.shape[0]
    if odom_jj.shape[0] != Eo:
        raise RuntimeError("odom_ii and odom_jj must have the same length")
    stride = 0
    if Eo > 0:
        _check(Tij, "Tij", torch.float32, dev)
        if Tij.dim() != 2 or Tij.shape[0] != Eo or Tij.shape[1] < 7:
            raise RuntimeError("Tij must be [num_odom_edges, 7 or 8]")
        stride = Tij.shape[1]
    sb = torch.as_tensor(s_bar).to(device=dev, dtype=torch.float64).contiguous()
    if sb.numel() != num_kf:
        raise RuntimeError("s_bar must hold one scale target per keyframe")
    p = GnParams(GN_RAYS, float(sigma_ray), float(sigma_dist), float(C_thresh), float(Q_thresh),
                 0, 0, 0, 0.0)
    lib = _lib.load()
    with torch.cuda.device(dev):
        dx = torch.zeros((max(num_kf - num_fix, 0), 7), dtype=torch.float32, device=dev)
        if max_iter <= 0 or num_kf - num_fix <= 0:
            return [dx]
        nbytes = lib.frodo_gn_odom_workspace_bytes(num_kf, E, N, Eo)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.frodo_gauss_newton_rays_odom(
            ctypes.byref(p), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj), _ptr(idx_ii2jj),
            _ptr(valid_match), _ptr(Q), _ptr(odom_ii), _ptr(odom_jj), _ptr(Tij) if Eo else None,
            int(stride), int(Eo), _ptr(sb), float(sigma_odom_t), float(sigma_odom_r),
            float(sigma_scale_prior), num_kf, N, E, int(num_fix), int(max_iter),
            float(delta_thresh), _ptr(dx), None, _ptr(ws), nbytes, _stream()))
    return [dx]


def make_params(mode, **kw):
    """GnParams from keyword names of the reference ops (used by dist.py / tests / bench)."""
    if mode == "rays":
        return GnParams(GN_RAYS, kw["sigma_ray"], kw["sigma_dist"], kw["C_thresh"], kw["Q_thresh"],
                        0, 0, 0, 0.0)
    if mode == "points":
        return GnParams(GN_POINTS, kw["sigma_point"], 0.0, kw["C_thresh"], kw["Q_thresh"], 0, 0, 0,
                        0.0)
    if mode == "calib":
        return GnParams(GN_CALIB, kw["sigma_pixel"], kw["sigma_depth"], kw["C_thresh"],
                        kw["Q_thresh"], kw["height"], kw["width"], kw["pixel_border"], kw["z_eps"])
    raise ValueError(mode)


def gn_system(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, num_fix):
    """One accumulation pass without the solve: returns (H [n,n] f64, g [n] f64,
    edge_blocks [E,35] f64, ii_row, jj_row).  Exposes the intermediate the reference
    keeps internal (Hs/gs, gn_kernels.cu:941-974) for parity tests and dist.py."""
    dev, num_kf, N, E = _gn_validate(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, K)
    lib = _lib.load()
    n = 7 * (num_kf - num_fix)
    with torch.cuda.device(dev):
        nbytes = lib.frodo_gn_workspace_bytes(num_kf, E, N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        Hg = torch.empty((n + 1) * n, dtype=torch.float64, device=dev)
        eb = torch.empty((E, 35), dtype=torch.float64, device=dev)
        ii_row = torch.empty(E, dtype=torch.int32, device=dev)
        jj_row = torch.empty(E, dtype=torch.int32, device=dev)
        st = _stream()
        _lib.check(lib.frodo_gn_workspace_init(_ptr(ws), nbytes, num_kf, E, N, st))
        _lib.check(lib.frodo_gn_edge_index(_ptr(ii), _ptr(jj), E, _ptr(ii_row), _ptr(jj_row), None,
                                           st))
        xc = None
        if N % 16 == 0 and E > 0:
            off = lib.frodo_gn_workspace_packed_offset(num_kf, E, N)
            xc = ctypes.c_void_p(ws.data_ptr() + off)
            _lib.check(lib.frodo_gn_pack_points(_ptr(Xs), _ptr(Cs), num_kf, N, xc, st))
        _lib.check(lib.frodo_gn_accumulate(
            ctypes.byref(params), _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(K), xc, _ptr(ii_row),
            _ptr(jj_row), _ptr(idx_ii2jj), _ptr(valid_match), _ptr(Q), num_kf, N, E, int(num_fix),
            _ptr(Hg), _ptr(eb), None, _ptr(ws), nbytes, st))
    return Hg[:n * n].view(n, n), Hg[n * n:], eb, ii_row, jj_row
