"""The allocation-free tracker entry point (frodo_track_pose_static: scratch in a static device
array inside the library, no counterpart in the reference) against the existing entry point
(frodo_track_pose, caller-provided workspace), through the C ABI, on two calls in a row.

Collected last on purpose (the file name sorts after every other test file).  Passes alone in a
fresh process and after the other tests: it only relies on the library."""
import ctypes

import numpy as np
import pytest
import torch

from frodo_slam_b200 import _lib, synth
from frodo_slam_b200._lib import GN_RAYS, TrackParams

pytestmark = pytest.mark.gpu

POSE_TOL = 1e-4  # the bar of tests/test_gpu_track.py (the two entry points are in fact bit-identical)


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _params():
    tp = synth.TRACK_PARAMS
    p = TrackParams()
    p.mode = GN_RAYS
    p.sigma_a, p.sigma_b = tp["sigma_ray"], tp["sigma_dist"]
    p.huber, p.max_iters = tp["huber"], tp["max_iters"]
    p.rel_error, p.delta_norm = tp["rel_error"], tp["delta_norm"]
    p.height = p.width = p.pixel_border = p.raw_calib = 0
    p.z_eps = 0.0
    return p


def _case(H, W, seed, dev):
    c = synth.tracking_case(H, W, seed=seed)
    t = {k: (v.to(dev).contiguous() if torch.is_tensor(v) else v) for k, v in c.items()}
    t["valid_u8"] = t["valid"].view(-1).to(torch.uint8).contiguous()
    t["Qk_f"] = t["Qk"].view(-1).contiguous()
    t["Tf"] = t["T_WCf"].view(-1).contiguous()
    t["Tk"] = t["T_WCk"].view(-1).contiguous()
    return t


def _outputs(p, dev):
    return (torch.zeros(8, device=dev), torch.zeros(8, device=dev),
            torch.full((2,), -1, dtype=torch.int32, device=dev),
            torch.zeros(max(p.max_iters, 1), dtype=torch.float64, device=dev))


def _run(lib, p, t, N, dev, static):
    out_f, out_rel, info, costs = _outputs(p, dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = (ctypes.byref(p), _ptr(t["Xf_all"]), _ptr(t["idx_f2k"]), _ptr(t["Xk"]), _ptr(t["Qk_f"]),
            _ptr(t["valid_u8"]), None, None, None, _ptr(t["Tf"]), _ptr(t["Tk"]), N, _ptr(out_f),
            _ptr(out_rel), _ptr(info), _ptr(costs))
    if static:
        _lib.check(lib.frodo_track_pose_static(*head, stream))
    else:
        nbytes = lib.frodo_track_workspace_bytes(N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.frodo_track_pose(*head, _ptr(ws), nbytes, stream))
    torch.cuda.synchronize()
    return out_f.cpu().numpy(), out_rel.cpu().numpy(), info.cpu().numpy(), costs.cpu().numpy()


def test_static_scratch_entry_matches_workspace_entry_twice():
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    assert lib.frodo_track_pose_static_max_points() == 512 * 384
    p = _params()
    # two calls in a row on the shared scratch: a large frame, then a smaller one with other data
    # (the second call sees the first one's leftovers in the array)
    for (H, W, seed) in ((96, 128, 1), (37, 53, 2)):
        with torch.cuda.device(dev):
            t = _case(H, W, seed, dev)
            N = H * W
            ref = _run(lib, p, t, N, dev, static=False)
            got = _run(lib, p, t, N, dev, static=True)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(dev)[0]  # driver-level: sees any cudaMalloc
            got2 = _run(lib, p, t, N, dev, static=True)  # and again, back to back
            free1 = torch.cuda.mem_get_info(dev)[0]
        assert free1 >= free0 - (64 << 10), "the static entry point allocated device memory"
        assert int(ref[2][1]) == 0 and int(ref[2][0]) >= 1
        for g in (got, got2):
            assert np.array_equal(g[2], ref[2]), "iterations / failure flag differ"
            assert np.abs(g[0] - ref[0]).max() <= POSE_TOL and np.abs(g[1] - ref[1]).max() <= POSE_TOL
            assert np.array_equal(g[0], ref[0]) and np.array_equal(g[1], ref[1]), "not bit-identical"
            it = int(ref[2][0])
            assert np.allclose(g[3][:it], ref[3][:it], rtol=1e-12, atol=0)


def test_static_scratch_entry_rejects_oversize():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    p = _params()
    N = lib.frodo_track_pose_static_max_points() + 1
    with torch.cuda.device(dev):
        x = torch.zeros(8, device=dev)
        big = torch.zeros(16, device=dev)  # never dereferenced: the size check comes first
        rc = lib.frodo_track_pose_static(
            ctypes.byref(p), _ptr(big), None, _ptr(big), _ptr(big), _ptr(big), None, None, None,
            _ptr(x), _ptr(x), N, _ptr(x), _ptr(x), None, None,
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 1  # FRODO_ERR_ARG
    assert b"exceeds the built-in scratch" in lib.frodo_last_error()
