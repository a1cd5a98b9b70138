"""The tracker's frame-to-keyframe pose solve on the B200.

Mirrors ``FrameTracker.opt_pose_ray_dist_sim3`` / ``opt_pose_calib_sim3``
(mast3r_slam/tracker.py:414-517; ``solve`` :397-412, geometry.py:17-104,
nonlinear_optimizer.py:4-33): same argument order and meaning, same return pair
``(T_WCf, T_CkCf)``, a failed Cholesky raises (the reference's
``torch.linalg.cholesky`` does, and ``track()`` catches it, tracker.py:312-314).
The reference runs ~15 PyTorch kernels and one host round trip per iteration;
here the whole solve is one persistent cooperative kernel behind
``frodo_track_pose`` (include/frodo_b200.h, csrc/track_pose.cu).

Poses may be ``lietorch.Sim3`` objects (anything with a ``.data`` tensor ``[1,8]``
and a constructor taking such a tensor) or plain ``[1,8]`` / ``[8]`` tensors
``[t, q_xyzw, s]``; the result has the type that was passed in.  No CPU fallback.
"""
import ctypes

import torch

from . import _lib
from ._lib import TrackParams, GN_RAYS, GN_CALIB

__all__ = ["track_pose", "opt_pose_ray_dist_sim3", "opt_pose_calib_sim3", "calib_measurements",
           "TRACK_DEFAULTS"]

# config/base.yaml:18-35 (`tracking`)
TRACK_DEFAULTS = dict(max_iters=50, rel_error=1e-3, delta_norm=1e-3, huber=1.345,
                      sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0, sigma_depth=10.0,
                      pixel_border=-10, depth_eps=1e-6)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None


def _pose_tensor(T, name, dev):
    d = T if isinstance(T, torch.Tensor) else getattr(T, "data", None)
    if not isinstance(d, torch.Tensor) or d.numel() != 8:
        raise RuntimeError(f"{name} must hold one Sim3 pose [t, q_xyzw, s] (8 values)")
    if not d.is_cuda or d.device != dev:
        raise RuntimeError(f"{name} must be a CUDA tensor on {dev} (frodo-slam_b200 has no CPU path)")
    if d.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32")
    return d.reshape(8).contiguous()


def _like(T, data):
    if isinstance(T, torch.Tensor):
        return data.reshape(T.shape)
    return type(T)(data.reshape(T.data.shape))


def _arr(t, name, dtype, n, cols, dev):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f"{name} must be a tensor")
    if not t.is_cuda or t.device != dev:
        raise RuntimeError(f"{name} must be a CUDA tensor on {dev} (frodo-slam_b200 has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must have dtype {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if n is not None and t.numel() != n * cols:
        raise RuntimeError(f"{name} must hold {n} x {cols} values, got {tuple(t.shape)}")
    return t


def track_pose(mode, Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k=None, valid_meas_k=None, K=None,
               img_size=None, idx_f2k=None, cfg=None):
    """Asynchronous core: returns dict(T_WCf [8], T_CkCf [8], info int32[2] = (iterations,
    failed), costs f64 [max_iters]) of device tensors; nothing is read back."""
    c = dict(TRACK_DEFAULTS)
    if cfg:
        c.update({k: cfg[k] for k in c if k in cfg})
    if mode not in ("rays", "calib"):
        raise ValueError(f"mode must be 'rays' or 'calib', got {mode!r}")
    if not isinstance(Qk, torch.Tensor) or not Qk.is_cuda:
        raise RuntimeError("Qk must be a CUDA tensor (frodo-slam_b200 has no CPU path)")
    dev = Qk.device
    N = Qk.numel()
    Qk = _arr(Qk, "Qk", torch.float32, None, 1, dev)
    valid = _arr(valid, "valid", torch.bool, N, 1, dev)
    if idx_f2k is not None:
        idx_f2k = _arr(idx_f2k, "idx_f2k", torch.int64, N, 1, dev)
        Xf = _arr(Xf, "Xf", torch.float32, None, 3, dev)
        if Xf.numel() % 3:
            raise RuntimeError("Xf must be [*,3]")
    else:
        Xf = _arr(Xf, "Xf", torch.float32, N, 3, dev)
    prm = TrackParams()
    prm.mode = GN_RAYS if mode == "rays" else GN_CALIB
    prm.huber = float(c["huber"])
    prm.max_iters = int(c["max_iters"])
    prm.rel_error = float(c["rel_error"])
    prm.delta_norm = float(c["delta_norm"])
    prm.height = prm.width = prm.pixel_border = prm.raw_calib = 0
    prm.z_eps = 0.0
    if mode == "rays":
        prm.sigma_a, prm.sigma_b = float(c["sigma_ray"]), float(c["sigma_dist"])
        Xk = _arr(Xk, "Xk", torch.float32, N, 3, dev)
    else:
        prm.sigma_a, prm.sigma_b = float(c["sigma_pixel"]), float(c["sigma_depth"])
        if meas_k is None:  # raw pointmaps: the kernel builds the measurement (raw_calib)
            prm.raw_calib = 1
            Xk = _arr(Xk, "Xk", torch.float32, N, 3, dev)
            if int(img_size[0]) * int(img_size[1]) != N:
                raise RuntimeError("raw calibrated tracking needs img_size[0] * img_size[1] == N")
        else:
            meas_k = _arr(meas_k, "meas_k", torch.float32, N, 3, dev)
            valid_meas_k = _arr(valid_meas_k, "valid_meas_k", torch.bool, N, 1, dev)
        K = _arr(K, "K", torch.float32, 3, 3, dev)
        prm.height, prm.width = int(img_size[0]), int(img_size[1])
        prm.pixel_border = int(c["pixel_border"])
        prm.z_eps = float(c["depth_eps"])
    Tf = _pose_tensor(T_WCf, "T_WCf", dev)
    Tk = _pose_tensor(T_WCk, "T_WCk", dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        # one allocation for the outputs (the kernel writes poses and info unconditionally and
        # costs[:iterations]), one for the workspace; nothing needs zeroing by the caller
        outs = torch.empty(max(prm.max_iters, 1) + 9, dtype=torch.float64, device=dev)
        costs = outs[9:]
        out_f, out_rel = outs[0:4].view(torch.float32), outs[4:8].view(torch.float32)
        info = outs[8:9].view(torch.int32)
        nbytes = lib.frodo_track_workspace_bytes(N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.frodo_track_pose(
            ctypes.byref(prm), _ptr(Xf), _ptr(idx_f2k),
            _ptr(Xk) if (mode == "rays" or prm.raw_calib) else None,
            _ptr(Qk), _ptr(valid), _ptr(meas_k) if (mode == "calib" and not prm.raw_calib) else None,
            _ptr(valid_meas_k) if (mode == "calib" and not prm.raw_calib) else None, _ptr(K) if mode == "calib" else None,
            _ptr(Tf), _ptr(Tk), N, _ptr(out_f), _ptr(out_rel), _ptr(info), _ptr(costs), _ptr(ws),
            nbytes, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return dict(T_WCf=out_f, T_CkCf=out_rel, info=info, costs=costs)


def _finish(r, T_WCf):
    iters, failed = r["info"].tolist()  # the one host read of the solve
    if failed:
        raise RuntimeError("track_pose: the 7x7 normal matrix is not positive definite "
                           "(torch.linalg.cholesky fails at tracker.py:407 in the reference)")
    return _like(T_WCf, r["T_WCf"]), _like(T_WCf, r["T_CkCf"])


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, idx_f2k=None):
    """tracker.py:414-465.  The reference is handed arrays already compacted by ``valid``
    (:293-295) and ignores ``valid``; both conventions work here: if the arrays have fewer rows
    than ``valid`` they are taken as compacted, otherwise ``valid`` masks them (no compaction
    pass, and ``idx_f2k`` lets the kernel gather ``Xf`` itself)."""
    n = Qk.numel()
    if valid is None or valid.numel() != n:
        valid = torch.ones(n, dtype=torch.bool, device=Qk.device)
    r = track_pose("rays", Xf, Xk, T_WCf, T_WCk, Qk, valid, idx_f2k=idx_f2k, cfg=cfg)
    return _finish(r, T_WCf)


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                        cfg=None, idx_f2k=None):
    """tracker.py:467-517 (``Xk`` is unused there too: the measurement is ``meas_k``).
    With ``meas_k=None`` the RAW pointmaps go in (``Xf`` the frame's own ``[H*W,3]`` with
    ``idx_f2k``, ``Xk`` the keyframe's) and the kernel's prologue does the calibrated branch of
    ``get_points_poses`` (tracker.py:383-393) itself."""
    r = track_pose("calib", Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                   idx_f2k=idx_f2k, cfg=cfg)
    return _finish(r, T_WCf)


def calib_measurements(Xf, Xk, K, img_size, depth_eps=1e-6):
    """The calibrated branch of ``FrameTracker.get_points_poses`` (tracker.py:383-393 with
    geometry.py:37-42,107-124): torch glue on the tensors' device, kept as in the reference
    (it is three elementwise passes per frame, not part of the iteration).  Returns the two
    pointmaps re-projected onto their pixel rays, ``meas_k`` and ``valid_meas_k``."""
    h, w = img_size
    dev, dt = Xf.device, Xf.dtype
    u, v = torch.meshgrid(torch.arange(w, device=dev), torch.arange(h, device=dev), indexing="xy")
    uv = torch.stack((u, v), dim=-1).to(dt).view(-1, 2)

    def constrain(X):
        z = X.reshape(-1, 3)[:, 2:3]
        d = torch.stack(((uv[:, 0] - K[0, 2]) / K[0, 0], (uv[:, 1] - K[1, 2]) / K[1, 1],
                         torch.ones_like(uv[:, 0])), dim=-1)
        return (z * d).contiguous()

    Xf_c, Xk_c = constrain(Xf), constrain(Xk)
    meas_k = torch.cat((uv, torch.log(Xk_c[:, 2:3])), dim=-1)
    valid_meas_k = Xk_c[:, 2:3] > depth_eps
    meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
    return Xf_c, Xk_c, meas_k.contiguous(), valid_meas_k.contiguous()
