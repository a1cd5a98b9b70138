"""GPU parity of the tracker's pose solve (mast3r_slam/tracker.py:397-517)
through the C ABI (frodo_track_pose via frodo_slam_b200.tracking).

Bars: iteration count and stop decision equal to the oracle's / the reference's; per-iteration
costs within 1e-5 relative; poses within 1e-4 (the sums over the points are re-associated:
fp32 per thread -> warp / CTA tree -> fp64 across CTAs, against torch's fp32 GEMM); run-to-run
and gather-vs-pregathered results bit-identical."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle
from frodo_slam_b200 import synth

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_track.npz")
TP = dict(synth.TRACK_PARAMS)


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def ft():
    from frodo_slam_b200 import tracking
    return tracking


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _case(H, W, seed, noise=0.01, **kw):
    c = synth.tracking_case(H, W, seed=seed, pose_noise=noise, **kw)
    return {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in c.items()}


def _oracle(mode, c, cfg=TP, valid=None, max_iters=None):
    valid = c["valid"] if valid is None else valid
    mi = cfg["max_iters"] if max_iters is None else max_iters
    if mode == "rays":
        return oracle.track_pose_rays(c["Xf_all"], c["Xk"], c["T_WCf"], c["T_WCk"], c["Qk"], valid,
                                      cfg["sigma_ray"], cfg["sigma_dist"], cfg["huber"], mi,
                                      cfg["rel_error"], cfg["delta_norm"], idx_f2k=c["idx_f2k"])
    hw = (c["H"], c["W"])
    Xf2, Xk2, meas, vmk = oracle.track_meas_calib(c["Xf_all"], c["Xk"], c["K"], hw,
                                                  cfg["depth_eps"])
    return oracle.track_pose_calib(Xf2, Xk2, c["T_WCf"], c["T_WCk"], c["Qk"], valid, meas, vmk,
                                   c["K"], hw, cfg["sigma_pixel"], cfg["sigma_depth"],
                                   cfg["pixel_border"], cfg["depth_eps"], cfg["huber"], mi,
                                   cfg["rel_error"], cfg["delta_norm"], idx_f2k=c["idx_f2k"])


def _gpu(ft, mode, c, dev, cfg=TP, valid=None, gather=True, max_iters=None):
    cfg = dict(cfg)
    if max_iters is not None:
        cfg["max_iters"] = max_iters
    valid = _t(c["valid"] if valid is None else valid, dev)
    Xf_all, idx = _t(c["Xf_all"], dev), _t(c["idx_f2k"], dev)
    if mode == "rays":
        Xf = Xf_all if gather else Xf_all[idx].contiguous()
        r = ft.track_pose("rays", Xf, _t(c["Xk"], dev), _t(c["T_WCf"], dev), _t(c["T_WCk"], dev),
                          _t(c["Qk"], dev), valid, idx_f2k=idx if gather else None, cfg=cfg)
    else:
        hw = (c["H"], c["W"])
        Xf_c, Xk_c, meas, vmk = ft.calib_measurements(Xf_all, _t(c["Xk"], dev), _t(c["K"], dev),
                                                      hw, cfg["depth_eps"])
        Xf = Xf_c if gather else Xf_c[idx].contiguous()
        r = ft.track_pose("calib", Xf, Xk_c, _t(c["T_WCf"], dev), _t(c["T_WCk"], dev),
                          _t(c["Qk"], dev), valid, meas, vmk, _t(c["K"], dev), hw,
                          idx_f2k=idx if gather else None, cfg=cfg)
    torch.cuda.synchronize()
    iters, failed = r["info"].tolist()
    return dict(T_WCf=r["T_WCf"].cpu().numpy(), T_CkCf=r["T_CkCf"].cpu().numpy(), iters=iters,
                status=failed, costs=r["costs"].cpu().numpy()[:iters])


def _agree(a, b, pose_tol=1e-4):
    assert a["status"] == b["status"] == 0
    assert a["iters"] == b["iters"], (a["iters"], b["iters"], a["costs"], b["costs"])
    assert np.abs(a["costs"] / b["costs"] - 1).max() < 1e-5
    assert np.abs(a["T_WCf"] - b["T_WCf"]).max() < pose_tol
    assert np.abs(a["T_CkCf"] - b["T_CkCf"]).max() < pose_tol


@pytest.mark.parametrize("mode", ["rays", "calib"])
@pytest.mark.parametrize("tag", ["a", "b"])
def test_golden_reference_tracker_replayed(ft, dev, mode, tag):
    """The reference's own tracker.py outputs (recorded with a stub lietorch)."""
    g = np.load(GOLD)
    H, W = (int(x) for x in g[f"{tag}_HW"])
    c = {k: g[f"{tag}_{k}"] for k in ("Xf_all", "idx_f2k", "Xk", "Qk", "valid", "T_WCf", "T_WCk", "K")}
    c.update(H=H, W=W)
    r = _gpu(ft, mode, c, dev)
    ref = dict(T_WCf=g[f"{tag}_{mode}_T_WCf"].reshape(8), T_CkCf=g[f"{tag}_{mode}_T_CkCf"].reshape(8),
               costs=g[f"{tag}_{mode}_costs"], iters=len(g[f"{tag}_{mode}_costs"]), status=0)
    _agree(r, ref)


@pytest.mark.parametrize("mode", ["rays", "calib"])
@pytest.mark.parametrize("H,W,seed,noise", [(48, 64, 1, 0.01), (37, 53, 2, 0.02), (9, 13, 3, 0.005),
                                            (120, 160, 5, 0.03)])
def test_vs_oracle(ft, dev, mode, H, W, seed, noise):
    c = _case(H, W, seed, noise)
    _agree(_gpu(ft, mode, c, dev), _oracle(mode, c))


@pytest.mark.parametrize("mode", ["rays", "calib"])
def test_full_size_512x384(ft, dev, mode):
    c = _case(384, 512, 7, 0.01)
    _agree(_gpu(ft, mode, c, dev), _oracle(mode, c))


def test_points_beyond_shared_memory_capacity(ft, dev):
    """> 148 x 6400 points: the tail of every CTA's slice iterates from the global record buffer."""
    c = _case(1000, 1024, 4, 0.01)
    _agree(_gpu(ft, "rays", c, dev), _oracle("rays", c))


@pytest.mark.parametrize("mode", ["rays", "calib"])
def test_fixed_iteration_counts(ft, dev, mode):
    c = _case(48, 64, 6, 0.02)
    strict = dict(TP, rel_error=0.0, delta_norm=0.0)
    for n in (1, 3):
        a, b = _gpu(ft, mode, c, dev, cfg=strict, max_iters=n), _oracle(mode, c, strict, max_iters=n)
        assert a["iters"] == n
        _agree(a, b)


@pytest.mark.parametrize("mode", ["rays", "calib"])
def test_gather_inside_kernel_equals_pregathered_bitwise(ft, dev, mode):
    c = _case(48, 64, 8, 0.01)
    a, b = _gpu(ft, mode, c, dev, gather=True), _gpu(ft, mode, c, dev, gather=False)
    assert a["iters"] == b["iters"]
    assert np.array_equal(a["T_WCf"], b["T_WCf"]) and np.array_equal(a["costs"], b["costs"])


def test_run_to_run_bitwise(ft, dev):
    c = _case(96, 128, 9, 0.02)
    a = _gpu(ft, "rays", c, dev)
    for _ in range(3):
        b = _gpu(ft, "rays", c, dev)
        assert np.array_equal(a["T_WCf"], b["T_WCf"]) and np.array_equal(a["costs"], b["costs"])


def test_reference_signature_with_compacted_arrays(ft, dev):
    """tracker.py:291-298 hands opt_pose_ray_dist_sim3 arrays compacted by valid_opt."""
    c = _case(48, 64, 10, 0.01)
    m = c["valid"][:, 0]
    Xf = c["Xf_all"][c["idx_f2k"]]

    class Pose:  # what a lietorch.Sim3 looks like to the operator
        def __init__(self, data):
            self.data = data

    T_f, T_k = Pose(_t(c["T_WCf"], dev)), Pose(_t(c["T_WCk"], dev))
    Tn, Trel = ft.opt_pose_ray_dist_sim3(_t(Xf[m], dev), _t(c["Xk"][m], dev), T_f, T_k,
                                         _t(c["Qk"][m], dev), _t(c["valid"], dev), cfg=TP)
    assert isinstance(Tn, Pose) and Tn.data.shape == (1, 8)
    ref = _oracle("rays", c)
    assert np.abs(Tn.data.cpu().numpy().reshape(8) - ref["T_WCf"]).max() < 1e-4
    assert np.abs(Trel.data.cpu().numpy().reshape(8) - ref["T_CkCf"]).max() < 1e-4
    # masked, full-size call gives the same answer
    Tn2, _ = ft.opt_pose_ray_dist_sim3(_t(Xf, dev), _t(c["Xk"], dev), T_f.data, T_k.data,
                                       _t(c["Qk"], dev), _t(c["valid"], dev), cfg=TP)
    assert torch.is_tensor(Tn2) and np.abs(Tn2.cpu().numpy().reshape(8) - ref["T_WCf"]).max() < 1e-4


def test_calib_signature(ft, dev):
    c = _case(48, 64, 11, 0.01)
    hw = (48, 64)
    Xf_c, Xk_c, meas, vmk = ft.calib_measurements(_t(c["Xf_all"], dev), _t(c["Xk"], dev),
                                                  _t(c["K"], dev), hw, TP["depth_eps"])
    o = oracle.track_meas_calib(c["Xf_all"], c["Xk"], c["K"], hw, TP["depth_eps"])
    assert np.abs(Xf_c.cpu().numpy() - o[0]).max() < 1e-6
    assert np.abs(meas.cpu().numpy() - o[2]).max() < 1e-6
    idx = _t(c["idx_f2k"], dev)
    Tn, Trel = ft.opt_pose_calib_sim3(Xf_c[idx].contiguous(), Xk_c, _t(c["T_WCf"], dev),
                                      _t(c["T_WCk"], dev), _t(c["Qk"], dev), _t(c["valid"], dev),
                                      meas, vmk, _t(c["K"], dev), hw, cfg=TP)
    ref = _oracle("calib", c)
    assert np.abs(Tn.cpu().numpy().reshape(8) - ref["T_WCf"]).max() < 1e-4


def test_no_valid_point_raises_like_cholesky(ft, dev):
    c = _case(16, 24, 2)
    none = np.zeros_like(c["valid"])
    r = _gpu(ft, "rays", c, dev, valid=none)
    assert r["status"] == 1 and r["iters"] == 0
    with pytest.raises(RuntimeError):
        ft.opt_pose_ray_dist_sim3(_t(c["Xf_all"][c["idx_f2k"]], dev), _t(c["Xk"], dev),
                                  _t(c["T_WCf"], dev), _t(c["T_WCk"], dev), _t(c["Qk"], dev),
                                  _t(none, dev), cfg=TP)


def test_invalid_points_may_hold_garbage(ft, dev):
    """Zero-information points never reach the sums, whatever they contain (the reference
    compacts them away, tracker.py:293-295)."""
    c = _case(48, 64, 12, 0.01)
    ref = _oracle("rays", c)
    bad = ~c["valid"][:, 0]
    c2 = dict(c)
    c2["Xk"] = c["Xk"].copy()
    c2["Xk"][bad] = np.nan
    c2["Qk"] = c["Qk"].copy()
    c2["Qk"][bad] = -1.0
    _agree(_gpu(ft, "rays", c2, dev), ref)


def test_zero_iterations_and_argument_errors(ft, dev):
    c = _case(16, 24, 2)
    r = _gpu(ft, "rays", c, dev, max_iters=0)
    assert r["iters"] == 0 and np.abs(r["T_WCf"] - c["T_WCf"].reshape(8)).max() < 1e-6
    with pytest.raises(RuntimeError):  # CPU tensors: no CPU path
        ft.track_pose("rays", torch.from_numpy(c["Xf"]), torch.from_numpy(c["Xk"]),
                      torch.from_numpy(c["T_WCf"]), torch.from_numpy(c["T_WCk"]),
                      torch.from_numpy(c["Qk"]), torch.from_numpy(c["valid"]))
    with pytest.raises(RuntimeError):  # wrong dtype
        ft.track_pose("rays", _t(c["Xf"], dev).double(), _t(c["Xk"], dev), _t(c["T_WCf"], dev),
                      _t(c["T_WCk"], dev), _t(c["Qk"], dev), _t(c["valid"], dev))
    with pytest.raises(RuntimeError):  # non-contiguous
        strided = _t(np.zeros((c["Xk"].shape[0], 6), np.float32), dev)[:, ::2]
        ft.track_pose("rays", _t(c["Xf"], dev), strided, _t(c["T_WCf"], dev), _t(c["T_WCk"], dev), _t(c["Qk"], dev), _t(c["valid"], dev))


def test_calib_from_raw_pointmaps_equals_torch_glue(ft, dev):
    """raw_calib: the prologue does constrain_points_to_ray + meas_k (tracker.py:383-393) itself."""
    for (H, W, seed) in ((48, 64, 11), (37, 53, 12)):
        c = _case(H, W, seed, 0.01)
        a = _gpu(ft, "calib", c, dev)  # torch glue (calib_measurements) + kernel
        r = ft.track_pose("calib", _t(c["Xf_all"], dev), _t(c["Xk"], dev), _t(c["T_WCf"], dev),
                          _t(c["T_WCk"], dev), _t(c["Qk"], dev), _t(c["valid"], dev), None, None,
                          _t(c["K"], dev), (H, W), idx_f2k=_t(c["idx_f2k"], dev), cfg=TP)
        it, failed = r["info"].tolist()
        assert failed == 0 and it == a["iters"]
        assert np.abs(r["T_WCf"].cpu().numpy() - a["T_WCf"]).max() < 1e-6
        assert np.abs(r["costs"].cpu().numpy()[:it] / a["costs"] - 1).max() < 1e-6
        _agree(dict(T_WCf=r["T_WCf"].cpu().numpy(), T_CkCf=r["T_CkCf"].cpu().numpy(), iters=it,
                    status=failed, costs=r["costs"].cpu().numpy()[:it]), _oracle("calib", c))
