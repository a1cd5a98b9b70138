"""The reference's ``mast3r_slam/matching.py`` on top of libfrodo_b200.

Same function names, arguments and return conventions as the reference module
(matching.py:8-95), so ``tracker.py`` / ``global_opt.py`` can import ``match`` from here:

    match(X11, X21, D11, D21, idx_1_to_2_init=None) -> (idx_1_to_2 [b,hw] i64, valid [b,hw,1] bool)

The reference runs ~20 PyTorch kernels around its two CUDA kernels (F.normalize, two depthwise
convolutions with reflect padding, concatenations, index conversions, a gather + norm, two
``.half()`` casts, thresholds).  Here ``match`` is ONE C-ABI call (``frodo_match``) that queues
seven kernels on torch's current stream: prep, iter_proj, occlusion test, two fp32->fp16
conversions, refine_matches, finish (csrc/match_glue.cu, iter_proj.cu, refine_matches.cu).
``prep_for_iter_proj`` and ``match_iterative_proj`` are also exposed step by step.

The matching section of the reference's config (config/base.yaml:9-16) is the ``cfg`` argument;
``integration/mast3r_slam/matching.py`` passes ``config["matching"]``.  No CPU path.
"""
import ctypes

import torch

from . import _lib
from ._lib import MatchParams
from .backends import _check, _check_contiguous, _ptr, _stream, iter_proj, refine_matches

__all__ = ["match", "match_iterative_proj", "prep_for_iter_proj", "pixel_to_lin", "lin_to_pixel",
           "match_quality", "DEFAULT_CFG"]

# config/base.yaml:9-16
DEFAULT_CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=1e-1,
                   radius=3, dilation_max=5, ft_matching_threshold=0.8)


def pixel_to_lin(p1, w):
    """matching.py:13-15."""
    return p1[..., 0] + (w * p1[..., 1])


def lin_to_pixel(idx_1_to_2, w):
    """matching.py:18-22."""
    u = idx_1_to_2 % w
    v = idx_1_to_2 // w
    return torch.stack((u, v), dim=-1)


def _params(cfg):
    c = dict(DEFAULT_CFG)
    if cfg is not None:
        c.update({k: cfg[k] for k in DEFAULT_CFG if k in cfg})
    return MatchParams(int(c["max_iter"]), float(c["lambda_init"]), float(c["convergence_thresh"]),
                       float(c["dist_thresh"]), int(c["radius"]), int(c["dilation_max"]),
                       float(c["ft_matching_threshold"])), c


def _check_maps(X11, X21, idx_1_to_2_init):
    _check_contiguous(X11=X11, X21=X21)
    _check(X11, "X11", torch.float32)
    dev = X11.device
    _check(X21, "X21", torch.float32, dev)
    if X11.dim() != 4 or X11.shape[3] != 3 or X21.shape != X11.shape:
        raise RuntimeError("X11 and X21 must both be [b,h,w,3]")
    b, h, w, _ = X11.shape
    if idx_1_to_2_init is not None:
        _check_contiguous(idx_1_to_2_init=idx_1_to_2_init)
        _check(idx_1_to_2_init, "idx_1_to_2_init", torch.int64, dev)
        if idx_1_to_2_init.shape != (b, h * w):
            raise RuntimeError("idx_1_to_2_init must be [b,h*w]")
    return b, h, w, dev


def prep_for_iter_proj(X11, X21, idx_1_to_2_init=None):
    """matching.py:25-49.  Returns (rays_with_grad_img [b,h,w,9], pts3d_norm [b,hw,3],
    p_init [b,hw,2] f32)."""
    b, h, w, dev = _check_maps(X11, X21, idx_1_to_2_init)
    lib = _lib.load()
    with torch.cuda.device(dev):
        rays = torch.empty((b, h, w, 9), dtype=torch.float32, device=dev)
        pts = torch.empty((b, h * w, 3), dtype=torch.float32, device=dev)
        p_init = torch.empty((b, h * w, 2), dtype=torch.float32, device=dev)
        _lib.check(lib.frodo_match_prep(_ptr(X11), _ptr(X21), _ptr(idx_1_to_2_init), _ptr(rays),
                                        _ptr(pts), _ptr(p_init), b, h, w, _stream()))
    return rays, pts, p_init


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None):
    """matching.py:52-95, one C-ABI call per stage (the operators stay individually usable and
    testable); ``match`` below is the single-call form."""
    prm, c = _params(cfg)
    b, h, w, dev = _check_maps(X11, X21, idx_1_to_2_init)
    rays, pts, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p, conv = iter_proj(rays, pts, p_init, c["max_iter"], c["lambda_init"],
                        c["convergence_thresh"])
    lib = _lib.load()
    with torch.cuda.device(dev):
        p1 = torch.empty((b, h * w, 2), dtype=torch.int64, device=dev)
        valid = torch.empty((b, h * w), dtype=torch.bool, device=dev)
        _lib.check(lib.frodo_match_occlusion(_ptr(X11), _ptr(X21), _ptr(p), _ptr(conv),
                                             float(c["dist_thresh"]), _ptr(p1), _ptr(valid), b, h,
                                             w, _stream()))
        halves = []
        for name, D in (("D11", D11), ("D21", D21)):
            _check_contiguous(**{name: D})
            if D.dtype == torch.float16:
                halves.append(D)
                continue
            _check(D, name, torch.float32, dev)
            Dh = torch.empty(D.shape, dtype=torch.float16, device=dev)
            _lib.check(lib.frodo_convert_f32_f16(_ptr(D), _ptr(Dh), D.numel(), _stream()))
            halves.append(Dh)
        p1, scores = refine_matches(halves[0], halves[1].view(b, h * w, -1), p1, c["radius"],
                                    c["dilation_max"])
        idx = torch.empty((b, h * w), dtype=torch.int64, device=dev)
        valid_out = torch.empty((b, h * w), dtype=torch.bool, device=dev)
        _lib.check(lib.frodo_match_finish(_ptr(p1), _ptr(scores), 1, _ptr(valid),
                                          float(c["ft_matching_threshold"]), _ptr(idx),
                                          _ptr(valid_out), b, h * w, w, _stream()))
    return idx, valid_out.unsqueeze(-1)


def match(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None):
    """matching.py:8-10.  X11, X21 [b,h,w,3] f32; D11, D21 [b,h,w,F] f32 (the reference casts
    them to fp16 itself, matching.py:79-80).  Returns (idx_1_to_2 [b,hw] i64,
    valid_match2 [b,hw,1] bool)."""
    prm, _ = _params(cfg)
    b, h, w, dev = _check_maps(X11, X21, idx_1_to_2_init)
    _check_contiguous(D11=D11, D21=D21)
    _check(D11, "D11", torch.float32, dev)
    _check(D21, "D21", torch.float32, dev)
    if D11.dim() != 4 or D11.shape[:3] != (b, h, w) or D21.numel() != D11.numel():
        raise RuntimeError("D11 must be [b,h,w,F] and D21 hold the same number of elements")
    F = D11.shape[3]
    lib = _lib.load()
    with torch.cuda.device(dev):
        nbytes = lib.frodo_match_workspace_bytes(b, h, w, F)
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
        off = (-ws.data_ptr()) % 256
        idx = torch.empty((b, h * w), dtype=torch.int64, device=dev)
        valid = torch.empty((b, h * w), dtype=torch.bool, device=dev)
        _lib.check(lib.frodo_match(ctypes.byref(prm), _ptr(X11), _ptr(X21), _ptr(D11), _ptr(D21),
                                   _ptr(idx_1_to_2_init), _ptr(idx), _ptr(valid), b, h, w, F,
                                   ctypes.c_void_p(ws.data_ptr() + off), nbytes, _stream()))
    return idx, valid.unsqueeze(-1)


def match_quality(Q_src, Q_dst, idx, valid_match, Q_conf, C_src=None, C_dst=None, C_conf=0.0):
    """What the callers derive from a match before using it, in one kernel
    (``frodo_match_quality``) instead of six torch kernels and two boolean sums:

    * the tracker (tracker.py:194-206): ``Qk = sqrt(Qff[idx_f2k] * Qkf)``, ``valid_opt =
      valid_match_k & (Cff[idx_f2k] > C_conf) & (Ckf > C_conf) & (Qk > Q_conf)``,
      ``valid_kf = valid_match_k & (Qk > Q_conf)``, ``match_frac = valid_opt.sum() / numel``;
    * ``FactorGraph.add_factors`` (global_opt.py:98-112): ``Qj = sqrt(Qii[b, idx_i2j] * Qji)``,
      ``valid_j = valid_match_j & (Qj > Q_conf)``, ``match_frac_j`` (no confidence test: leave
      ``C_src`` / ``C_dst`` unset).

    Q_src, C_src ``[B,N_src(,1)]`` (or unbatched), Q_dst, C_dst, valid_match ``[B,N(,1)]``, idx
    ``[B,N]`` int64.  Returns ``(Q, valid_opt, valid_q, counts)`` shaped like ``Q_dst`` /
    ``valid_match``, ``counts [B,2] int32 = (sum valid_opt, sum valid_q)`` on the device
    (``match_frac = counts[:, 0] / N``; nothing is read back here).  Bit-identical to torch."""
    batched = idx.dim() >= 2
    B = idx.shape[0] if batched else 1
    N = idx.numel() // max(B, 1)
    dev = idx.device
    named = dict(Q_src=Q_src, Q_dst=Q_dst, idx=idx, valid_match=valid_match)
    if C_src is not None:
        named["C_src"] = C_src
    if C_dst is not None:
        named["C_dst"] = C_dst
    _check_contiguous(**named)
    _check(idx, "idx", torch.int64)
    _check(valid_match, "valid_match", torch.bool, dev)
    for name in ("Q_src", "Q_dst", "C_src", "C_dst"):
        if name in named:
            _check(named[name], name, torch.float32, dev)
    N_src = Q_src.numel() // max(B, 1)
    if Q_dst.numel() != B * N or valid_match.numel() != B * N or Q_src.numel() != B * N_src or \
            (C_src is not None and C_src.numel() != B * N_src) or \
            (C_dst is not None and C_dst.numel() != B * N):
        raise RuntimeError("match_quality: Q_dst / C_dst / valid_match must be [B,N], Q_src / C_src "
                           "[B,N_src]")
    lib = _lib.load()
    with torch.cuda.device(dev):
        Q = torch.empty_like(Q_dst)
        valid_opt = torch.empty_like(valid_match)
        valid_q = torch.empty_like(valid_match)
        counts = torch.empty((B, 2), dtype=torch.int32, device=dev)
        _lib.check(lib.frodo_match_quality(
            _ptr(Q_src), _ptr(Q_dst), _ptr(C_src), _ptr(C_dst), _ptr(idx), _ptr(valid_match),
            float(Q_conf), float(C_conf), _ptr(Q), _ptr(valid_opt), _ptr(valid_q), _ptr(counts),
            B, N, N_src, _stream()))
    return Q, valid_opt, valid_q, counts

