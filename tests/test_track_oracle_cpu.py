"""CPU: the oracle's restatement of the tracker's pose solve (tracker.py:397-517)
against the reference's own Python (tests/golden/golden_track.npz, recorded by
tests/golden/make_golden_track.py with a stub lietorch), plus self-checks."""
import os

import numpy as np
import pytest

from oracle import oracle
from frodo_slam_b200 import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_track.npz")


@pytest.fixture(scope="module")
def gold():
    g = np.load(GOLD)
    cfg = {k[4:]: float(g[k]) for k in g.files if k.startswith("cfg_")}
    return g, cfg


def test_recorded_config_is_the_reference_default(gold):
    _, cfg = gold
    for k, v in synth.TRACK_PARAMS.items():
        assert cfg[k] == pytest.approx(v), k


def run_rays(g, t, cfg, **kw):
    return oracle.track_pose_rays(g[f"{t}_Xf_all"], g[f"{t}_Xk"], g[f"{t}_T_WCf"], g[f"{t}_T_WCk"],
                                  g[f"{t}_Qk"], g[f"{t}_valid"], cfg["sigma_ray"],
                                  cfg["sigma_dist"], cfg["huber"], int(cfg["max_iters"]),
                                  cfg["rel_error"], cfg["delta_norm"],
                                  idx_f2k=g[f"{t}_idx_f2k"], **kw)


def run_calib(g, t, cfg):
    H, W = (int(x) for x in g[f"{t}_HW"])
    Xf2, Xk2, meas, vmk = oracle.track_meas_calib(g[f"{t}_Xf_all"], g[f"{t}_Xk"], g[f"{t}_K"],
                                                  (H, W), cfg["depth_eps"])
    r = oracle.track_pose_calib(Xf2, Xk2, g[f"{t}_T_WCf"], g[f"{t}_T_WCk"], g[f"{t}_Qk"],
                                g[f"{t}_valid"], meas, vmk, g[f"{t}_K"], (H, W),
                                cfg["sigma_pixel"], cfg["sigma_depth"], int(cfg["pixel_border"]),
                                cfg["depth_eps"], cfg["huber"], int(cfg["max_iters"]),
                                cfg["rel_error"], cfg["delta_norm"], idx_f2k=g[f"{t}_idx_f2k"])
    return r, (Xf2, Xk2, meas, vmk)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_rays_oracle_matches_reference_tracker(gold, tag):
    g, cfg = gold
    r = run_rays(g, tag, cfg)
    ref_costs = g[f"{tag}_rays_costs"]
    assert r["status"] == 0 and r["iters"] == len(ref_costs)
    assert np.abs(r["costs"] / ref_costs - 1).max() < 1e-5
    # the ray residual leaves translation along the view weakly observable: poses to 1e-4
    assert np.abs(r["T_WCf"] - g[f"{tag}_rays_T_WCf"].reshape(8)).max() < 1e-4
    assert np.abs(r["T_CkCf"] - g[f"{tag}_rays_T_CkCf"].reshape(8)).max() < 1e-4


@pytest.mark.parametrize("tag", ["a", "b"])
def test_calib_oracle_matches_reference_tracker(gold, tag):
    g, cfg = gold
    r, (Xf2, Xk2, meas, vmk) = run_calib(g, tag, cfg)
    assert np.array_equal(Xf2, g[f"{tag}_calib_Xf"]) and np.array_equal(Xk2, g[f"{tag}_calib_Xk"])
    assert np.abs(meas - g[f"{tag}_calib_meas_k"]).max() < 2.4e-7
    assert np.array_equal(vmk, g[f"{tag}_calib_valid_meas_k"])
    ref_costs = g[f"{tag}_calib_costs"]
    assert r["status"] == 0 and r["iters"] == len(ref_costs)
    assert np.abs(r["costs"] / ref_costs - 1).max() < 1e-5
    assert np.abs(r["T_WCf"] - g[f"{tag}_calib_T_WCf"].reshape(8)).max() < 2e-5
    assert np.abs(r["T_CkCf"] - g[f"{tag}_calib_T_CkCf"].reshape(8)).max() < 2e-5


def test_sim3_group_ops_are_consistent():
    rng = np.random.default_rng(0)
    T = synth.random_poses(3, __import__("torch").Generator().manual_seed(3)).numpy().astype(np.float32)
    X = rng.normal(size=(50, 3)).astype(np.float32)
    A, B = T[0], T[1]
    AB = oracle.sim3_mul(A, B)
    assert np.abs(oracle.sim3_act(AB, X) - oracle.sim3_act(A, oracle.sim3_act(B, X))).max() < 1e-5
    I = oracle.sim3_mul(oracle.sim3_inv(A), A)
    assert np.abs(I - np.array([0, 0, 0, 0, 0, 0, 1, 1], np.float32)).max() < 1e-6
    assert np.abs(oracle.sim3_retr(A, np.zeros(7, np.float32)) - A).max() < 1e-7


def test_cost_decreases_and_pose_moves_towards_truth():
    c = synth.tracking_case(40, 56, seed=2, pose_noise=0.02)
    tp = synth.TRACK_PARAMS
    Xf2, Xk2, meas, vmk = oracle.track_meas_calib(c["Xf_all"].numpy(), c["Xk"].numpy(),
                                                  c["K"].numpy(), (40, 56), tp["depth_eps"])
    r = oracle.track_pose_calib(Xf2, Xk2, c["T_WCf"].numpy(), c["T_WCk"].numpy(), c["Qk"].numpy(),
                                c["valid"].numpy(), meas, vmk, c["K"].numpy(), (40, 56),
                                idx_f2k=c["idx_f2k"].numpy())
    assert r["iters"] >= 2 and np.all(np.diff(r["costs"]) < 0)
    err0 = np.abs(c["T_WCf"].numpy() - c["T_WCf_true"].numpy()).max()
    err1 = np.abs(r["T_WCf"] - c["T_WCf_true"].numpy().reshape(8)).max()
    assert err1 < err0


def test_no_valid_point_reports_cholesky_failure():
    c = synth.tracking_case(16, 24, seed=2)
    r = oracle.track_pose_rays(c["Xf"].numpy(), c["Xk"].numpy(), c["T_WCf"].numpy(),
                               c["T_WCk"].numpy(), c["Qk"].numpy(),
                               np.zeros_like(c["valid"].numpy()))
    assert r["status"] == 1 and r["iters"] == 0
