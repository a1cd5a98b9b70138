"""Pins the CPU oracle against golden vectors recorded from the UNMODIFIED reference
backend on a B200 (tests/golden/make_golden_ref.py -> tests/golden/golden_ref.npz)."""
import os

import numpy as np
import pytest

from helpers import GP, MP, rel_err
from oracle import oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_ref.npz")


@pytest.fixture(scope="module")
def gold():
    assert os.path.exists(GOLD), "golden fixture missing"
    return np.load(GOLD)


def test_iter_proj_oracle_equals_reference_bitwise(gold):
    from oracle import oracle
    p, c = oracle.iter_proj(gold["ip_rays"], gold["ip_pts"], gold["ip_p0"], MP["max_iter"],
                            MP["lambda_init"], MP["convergence_thresh"])
    assert np.array_equal(p.view(np.uint32), gold["ip_p"].view(np.uint32))
    assert np.array_equal(c, gold["ip_conv"])


def test_refine_f16_oracle_equals_reference_bitwise(gold):
    from oracle import oracle
    pn, sc = oracle.refine_matches(gold["rm_D11"], gold["rm_D21"], gold["rm_p1"], MP["radius"],
                                   MP["dilation_max"])
    assert np.array_equal(pn, gold["rm_pn"])
    assert np.array_equal(sc.view(np.uint16), gold["rm_sc"].view(np.uint16))


def test_refine_f32_oracle_equals_reference_bitwise(gold):
    from oracle import oracle
    pn, sc = oracle.refine_matches(gold["rm_D11"].astype(np.float32),
                                   gold["rm_D21"].astype(np.float32), gold["rm_p1"], 2, 3)
    assert np.array_equal(pn, gold["rm32_pn"])
    assert np.array_equal(sc.view(np.uint32), gold["rm32_sc"].view(np.uint32))


def _gn(gold, mode, num_fix, iters):
    from oracle import oracle
    H, W = (int(x) for x in gold["gn_HW"])
    kw = dict(rays=dict(sigma_a=GP["sigma_ray"], sigma_b=GP["sigma_dist"]),
              points=dict(sigma_a=GP["sigma_point"], sigma_b=0.0),
              calib=dict(sigma_a=GP["sigma_pixel"], sigma_b=GP["sigma_depth"], K=gold["gn_K"],
                         height=H, width=W, pixel_border=GP["pixel_border"],
                         z_eps=GP["depth_eps"]))[mode]
    return oracle.gauss_newton(mode, gold["gn_Twc"], gold["gn_Xs"], gold["gn_Cs"], gold["gn_ii"],
                               gold["gn_jj"], gold["gn_idx"].astype(np.int64), gold["gn_valid"],
                               gold["gn_Q"], C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"],
                               num_fix=num_fix, max_iter=iters, delta_thresh=0.0, **kw)


@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
@pytest.mark.parametrize("iters", [1, 3])
def test_gn_oracle_matches_reference(gold, mode, iters):
    r = _gn(gold, mode, 1, iters)
    assert r["fails"] == 0 and r["iters"] == iters
    assert rel_err(r["Twc"], gold[f"gn_{mode}_{iters}_Twc"]) < 1e-4
    if iters == 1:
        assert rel_err(r["dx"], gold[f"gn_{mode}_{iters}_dx"]) < 1e-4


def test_gn_oracle_num_fix_2(gold):
    r = _gn(gold, "rays", 2, 2)
    assert rel_err(r["Twc"], gold["gn_rays_fix2_Twc"]) < 1e-4
    assert np.array_equal(r["Twc"][:2], gold["gn_Twc"][:2])  # pinned poses untouched


# ------------------------------------------------ gauss_newton_rays_odom (golden_ref_odom.npz)

GOLD_ODOM = os.path.join(os.path.dirname(__file__), "golden", "golden_ref_odom.npz")


@pytest.mark.parametrize("num_fix,iters", [(1, 1), (1, 3), (2, 2)])
def test_gn_rays_odom_oracle_matches_reference(num_fix, iters):
    """Recorded from the unmodified reference by tests/golden/make_golden_ref_odom.py."""
    from helpers import OP
    from oracle import oracle
    gd = np.load(GOLD_ODOM)
    r = oracle.gauss_newton_rays_odom(
        gd["gn_Twc"], gd["gn_Xs"], gd["gn_Cs"], gd["gn_ii"], gd["gn_jj"],
        gd["gn_idx"].astype(np.int64), gd["gn_valid"], gd["gn_Q"], gd["odom_ii"], gd["odom_jj"],
        gd["Tij"], gd["s_bar"], OP["sigma_odom_t"], OP["sigma_odom_r"], OP["sigma_ray"],
        GP["sigma_dist"], OP["sigma_scale_prior"], GP["C_conf"], GP["Q_conf"], num_fix, iters, 0.0)
    assert r["fails"] == 0 and r["iters"] == iters
    assert rel_err(r["Twc"], gd[f"odom_fix{num_fix}_it{iters}_Twc"]) < 1e-4
    if iters == 1:
        assert rel_err(r["dx"], gd[f"odom_fix{num_fix}_it{iters}_dx"]) < 1e-4
    assert np.array_equal(r["Twc"][:num_fix], gd["gn_Twc"][:num_fix])


# ------------------------------------------------ match() pipeline (golden_match.npz)
GOLD_MATCH = os.path.join(os.path.dirname(__file__), "golden", "golden_match.npz")


def _match_cfg(g):
    c = {k[4:]: float(g[k]) for k in g.files if k.startswith("cfg_")}
    for k in ("max_iter", "radius", "dilation_max"):
        c[k] = int(c[k])
    return c


@pytest.mark.parametrize("tag", ["a", "b"])
def test_match_pipeline_oracle_vs_reference_python(tag):
    """Recorded from the reference's own matching.py / image.py (CPU torch) by
    tests/golden/make_golden_match.py.  Prologue: torch orders the sums differently, 2 ulp;
    the tail on the recorded intermediates and the whole pipeline: identical indices and masks."""
    g = np.load(GOLD_MATCH)
    c = _match_cfg(g)
    init = None if tag == "a" else g["idx_init"]
    rays, pts, p_init = oracle.prep_for_iter_proj(g["X11"], g["X21"], init)
    assert np.abs(rays - g[tag + "_rays_with_grad_img"]).max() <= 2.4e-7   # values are <= 1
    assert np.abs(pts - g[tag + "_pts3d_norm"]).max() <= 2.4e-7
    assert np.array_equal(p_init, g[tag + "_p_init"])
    p1, v = oracle.match_occlusion(g["X11"], g["X21"], g[tag + "_p_iter"], g[tag + "_conv"],
                                   c["dist_thresh"])
    idx, valid = oracle.match_finish(g[tag + "_p_refined"], g[tag + "_scores"], v,
                                     g["X11"].shape[2], c["ft_matching_threshold"])
    assert np.array_equal(idx, g[tag + "_idx_1_to_2"])
    assert np.array_equal(valid, g[tag + "_valid"])
    idx2, valid2 = oracle.match(g["X11"], g["X21"], g["D11"], g["D21"], init, **c)
    assert np.array_equal(idx2, g[tag + "_idx_1_to_2"])
    assert np.array_equal(valid2, g[tag + "_valid"])
    assert 0.2 < valid2.mean() < 0.95   # the fixture exercises both outcomes
