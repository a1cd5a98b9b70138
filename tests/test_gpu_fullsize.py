"""GPU parity tests, part 2: the committed golden fixture (outputs of the unmodified
reference recorded on a B200), ragged / empty inputs, BASELINE.json's full 512x384 size,
and size-independent properties (run-to-run determinism, edge-order invariance, the
split C-ABI calls equal the single call).  Everything goes through the C ABI."""
import os

import numpy as np
import pytest
import torch

from helpers import GP, MP, backend_gn, graph_inputs, matching_inputs, oracle_gn, rel_err

pytestmark = pytest.mark.gpu

GN_TOL = 1e-4
GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_ref.npz")


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def be():
    import mast3r_slam_backends
    return mast3r_slam_backends


@pytest.fixture(scope="module")
def gold():
    assert os.path.exists(GOLD), "golden fixture missing"
    return np.load(GOLD)


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


# ------------------------------------------------------- golden fixture (reference outputs)

def test_golden_iter_proj(be, dev, gold):
    p, c = be.iter_proj(_t(gold["ip_rays"], dev), _t(gold["ip_pts"], dev), _t(gold["ip_p0"], dev),
                        MP["max_iter"], MP["lambda_init"], MP["convergence_thresh"])
    assert np.array_equal(p.cpu().numpy().view(np.uint32), gold["ip_p"].view(np.uint32))
    assert np.array_equal(c.cpu().numpy(), gold["ip_conv"])


def test_golden_refine_matches(be, dev, gold):
    pn, sc = be.refine_matches(_t(gold["rm_D11"], dev), _t(gold["rm_D21"], dev),
                               _t(gold["rm_p1"], dev), MP["radius"], MP["dilation_max"])
    assert np.array_equal(pn.cpu().numpy(), gold["rm_pn"])
    assert np.array_equal(sc.cpu().numpy().view(np.uint16), gold["rm_sc"].view(np.uint16))
    pn, sc = be.refine_matches(_t(gold["rm_D11"].astype(np.float32), dev),
                               _t(gold["rm_D21"].astype(np.float32), dev), _t(gold["rm_p1"], dev),
                               2, 3)
    assert np.array_equal(pn.cpu().numpy(), gold["rm32_pn"])
    assert np.array_equal(sc.cpu().numpy().view(np.uint32), gold["rm32_sc"].view(np.uint32))


def _gold_graph(gold):
    H, W = (int(x) for x in gold["gn_HW"])
    return dict(Twc=torch.from_numpy(gold["gn_Twc"]), Xs=torch.from_numpy(gold["gn_Xs"]),
                Cs=torch.from_numpy(gold["gn_Cs"]), K=torch.from_numpy(gold["gn_K"]),
                ii=torch.from_numpy(gold["gn_ii"]), jj=torch.from_numpy(gold["gn_jj"]),
                idx_ii2jj=torch.from_numpy(gold["gn_idx"].astype(np.int64)),
                valid_match=torch.from_numpy(gold["gn_valid"]), Q=torch.from_numpy(gold["gn_Q"]),
                H=H, W=W)


@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
@pytest.mark.parametrize("iters", [1, 3])
def test_golden_gn(be, dev, gold, mode, iters):
    g = _gold_graph(gold)
    Twc, dx = backend_gn(be, mode, g, dev, 1, iters)
    assert rel_err(Twc, gold[f"gn_{mode}_{iters}_Twc"]) < GN_TOL
    if iters == 1:
        assert rel_err(dx, gold[f"gn_{mode}_{iters}_dx"]) < GN_TOL


def test_golden_gn_num_fix_2(be, dev, gold):
    g = _gold_graph(gold)
    Twc, _ = backend_gn(be, "rays", g, dev, 2, 2)
    assert rel_err(Twc, gold["gn_rays_fix2_Twc"]) < GN_TOL
    assert np.array_equal(Twc[:2], gold["gn_Twc"][:2])


# ----------------------------------------------------------------- ragged and empty inputs

@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
@pytest.mark.parametrize("H,W", [(37, 53), (33, 48)])  # N = 1961 (odd), 1584 (16 | N, not a tile)
def test_gn_ragged_sizes(be, dev, mode, H, W):
    g = graph_inputs("window8", H, W)
    Twc, dx = backend_gn(be, mode, g, dev, 1, 2)
    ref = oracle_gn(mode, g, 1, 2)
    assert ref["fails"] == 0
    assert rel_err(Twc, ref["Twc"]) < GN_TOL


def test_empty_inputs(be, dev):
    rays = torch.zeros(0, 8, 8, 9, device=dev)
    p, c = be.iter_proj(rays, torch.zeros(0, 64, 3, device=dev), torch.zeros(0, 64, 2, device=dev),
                        10, 1e-8, 1e-6)
    assert p.shape == (0, 64, 2) and c.shape == (0, 64)
    pn, sc = be.refine_matches(torch.zeros(0, 8, 8, 24, device=dev, dtype=torch.half),
                               torch.zeros(0, 64, 24, device=dev, dtype=torch.half),
                               torch.zeros(0, 64, 2, device=dev, dtype=torch.long), 3, 5)
    assert pn.shape == (0, 64, 2) and sc.shape == (0, 64)
    # a graph without edges: H = 0 -> factorisation fails -> dx = 0, poses untouched
    g = graph_inputs("window8", 16, 16)
    for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q"):
        g[k] = g[k][:0].contiguous()
    Twc, dx = backend_gn(be, "rays", g, dev, 1, 2)
    assert np.all(dx == 0) and np.array_equal(Twc, g["Twc"].numpy())


def test_gn_zero_iterations_and_all_fixed(be, dev):
    g = graph_inputs("pair", 16, 16)
    Twc, dx = backend_gn(be, "rays", g, dev, 1, 0)
    assert dx.shape == (1, 7) and np.all(dx == 0) and np.array_equal(Twc, g["Twc"].numpy())
    Twc, dx = backend_gn(be, "rays", g, dev, 2, 3)
    assert dx.shape == (0, 7) and np.array_equal(Twc, g["Twc"].numpy())


# ------------------------------------------------------------------ properties, any size

def _sharded(dev, g, mode="rays", num_fix=1, order=None):
    from frodo_slam_b200 import backends, dist as fdist
    prm = backends.make_params(mode, sigma_ray=GP["sigma_ray"], sigma_dist=GP["sigma_dist"],
                               C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"])
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    E, N = t["ii"].shape[0], t["Xs"].shape[1]
    if order is not None:
        o = torch.as_tensor(order, device=dev)
        for k in ("ii", "jj", "idx_ii2jj", "valid_match", "Q"):
            t[k] = t[k][o].contiguous()
    Twc = t["Twc"].clone()
    sg = fdist.ShardedGaussNewton(prm, Twc, t["Xs"], t["Cs"], None, t["ii"], t["jj"],
                                  t["idx_ii2jj"].view(E, N), t["valid_match"].view(E, N),
                                  t["Q"].view(E, N), 0, E, num_fix)
    return sg, Twc


def test_split_calls_equal_single_call_bitwise(be, dev):
    """accumulate / solve_retract issued separately (dist.py, world 1) == frodo_gauss_newton."""
    g = graph_inputs("window8", 48, 64)
    sg, Twc = _sharded(dev, g)
    dx = sg.solve(3, 0.0)
    torch.cuda.synchronize()
    Tb, db = backend_gn(be, "rays", g, dev, 1, 3)
    assert np.array_equal(Twc.cpu().numpy(), Tb) and np.array_equal(dx.cpu().numpy(), db)
    assert int(sg.flags[1]) == 3 and int(sg.flags[2]) == 0


@pytest.mark.parametrize("mode", ["rays", "calib"])
def test_gn_run_to_run_deterministic(be, dev, mode):
    g = graph_inputs("calib32", 64, 80)
    a = backend_gn(be, mode, g, dev, 1, 3)
    b = backend_gn(be, mode, g, dev, 1, 3)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_gn_edge_order_invariance(be, dev):
    """Permuting the edge list only re-orders fp64 sums of the assembly."""
    g = graph_inputs("window8", 48, 64)
    E = g["ii"].shape[0]
    perm = torch.randperm(E, generator=torch.Generator().manual_seed(4)).tolist()
    s0, T0 = _sharded(dev, g)
    s1, T1 = _sharded(dev, g, order=perm)
    s0.solve(2, 0.0)
    s1.solve(2, 0.0)
    torch.cuda.synchronize()
    assert rel_err(T1.cpu().numpy(), T0.cpu().numpy()) < 1e-6


# ----------------------------------------------------- BASELINE.json's full size (512x384)

def test_fullsize_matching_bit_exact(be, dev):
    from oracle import oracle
    H, W = 384, 512
    d, rays, pts, p0 = matching_inputs(1, H, W, seed=23)
    p, c = be.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), MP["max_iter"], MP["lambda_init"],
                        MP["convergence_thresh"])
    pr, cr = oracle.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), MP["max_iter"],
                              MP["lambda_init"], MP["convergence_thresh"])
    assert np.array_equal(p.cpu().numpy(), pr) and np.array_equal(c.cpu().numpy(), cr)
    D11 = d["D11"].half().contiguous()
    D21 = d["D21"].half().reshape(1, H * W, -1).contiguous()
    p1 = p.long()
    pn, sc = be.refine_matches(D11.to(dev), D21.to(dev), p1, MP["radius"], MP["dilation_max"])
    pnr, scr = oracle.refine_matches(D11.numpy(), D21.numpy(), p1.cpu().numpy(), MP["radius"],
                                     MP["dilation_max"])
    assert np.array_equal(pn.cpu().numpy(), pnr)
    assert np.array_equal(sc.cpu().numpy().view(np.uint16), scr.view(np.uint16))
    # property: the refined match never scores below the centre it started from
    assert pn.min() >= 0 and int(pn[..., 0].max()) < W and int(pn[..., 1].max()) < H


def test_fullsize_gn_system_and_solve_vs_oracle(be, dev):
    """configs[1] (8-keyframe window, E = 24) at 512x384: packed rays kernel, many tiles/CTA."""
    from frodo_slam_b200 import backends
    from oracle import oracle
    g = graph_inputs("window8", 384, 512)
    p = backends.make_params("rays", sigma_ray=GP["sigma_ray"], sigma_dist=GP["sigma_dist"],
                             C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"])
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    H, gv, eb, _, _ = backends.gn_system(p, t["Twc"], t["Xs"], t["Cs"], None, t["ii"], t["jj"],
                                         t["idx_ii2jj"], t["valid_match"], t["Q"], 1)
    torch.cuda.synchronize()
    ie, je, _ = oracle.edge_index(g["ii"].numpy(), g["jj"].numpy())
    Hs, gs = oracle.edge_blocks("rays", g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie, je,
                                g["idx_ii2jj"].numpy(), g["valid_match"].numpy(), g["Q"].numpy(),
                                GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"])
    Hr, gr = oracle.assemble(Hs, gs, ie, je, 8, 1, round_f32=False)
    assert rel_err(H.cpu().numpy(), Hr) < GN_TOL and rel_err(gv.cpu().numpy(), gr) < GN_TOL
    Twc, dx = backend_gn(be, "rays", g, dev, 1, 2)
    ref = oracle_gn("rays", g, 1, 2)
    assert rel_err(Twc, ref["Twc"]) < GN_TOL


# ------------------------------------------ gauss_newton_rays_odom

@pytest.mark.parametrize("cfg,num_fix,iters", [("window8", 1, 1), ("window8", 1, 3),
                                               ("window8", 2, 2), ("calib32", 1, 2)])
def test_gn_rays_odom_vs_oracle(be, dev, cfg, num_fix, iters):
    from helpers import backend_gn_odom, odom_inputs, oracle_gn_odom
    g = graph_inputs(cfg, 48, 64)
    od = odom_inputs(g)
    Twc, dx = backend_gn_odom(be, g, od, dev, num_fix, iters)
    ref = oracle_gn_odom(g, od, num_fix, iters)
    assert ref["fails"] == 0
    assert rel_err(Twc, ref["Twc"]) < GN_TOL
    if iters == 1:
        assert rel_err(dx, ref["dx"]) < GN_TOL
    assert np.array_equal(Twc[:num_fix], g["Twc"].numpy()[:num_fix])


def test_gn_rays_odom_without_odometry_equals_rays_plus_prior(be, dev):
    """No odometry edges and s_bar = 0 (every prior skipped) -> exactly gauss_newton_rays."""
    from helpers import OP, backend_gn_odom
    g = graph_inputs("window8", 48, 64)
    od = dict(odom_ii=torch.zeros(0, dtype=torch.long), odom_jj=torch.zeros(0, dtype=torch.long),
              Tij=torch.zeros(0, 8), s_bar=torch.zeros(8, dtype=torch.float64))
    Ta, da = backend_gn_odom(be, g, od, dev, 1, 2)
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    Twc = t["Twc"].clone()
    db = be.gauss_newton_rays(Twc, t["Xs"], t["Cs"], t["ii"], t["jj"], t["idx_ii2jj"],
                              t["valid_match"], t["Q"], OP["sigma_ray"], GP["sigma_dist"],
                              GP["C_conf"], GP["Q_conf"], 1, 2, 0.0)
    torch.cuda.synchronize()
    assert np.array_equal(Ta, Twc.cpu().numpy()) and np.array_equal(da, db[0].cpu().numpy())


def test_gn_rays_odom_vs_reference(be, dev, ref_backend, capfd):
    if ref_backend is None:
        pytest.skip("oracle/_ref not built")
    from helpers import backend_gn_odom, odom_inputs
    g = graph_inputs("window8", 48, 64)
    od = odom_inputs(g)
    for num_fix, iters in ((1, 1), (1, 3), (2, 2)):
        Ta, da = backend_gn_odom(be, g, od, dev, num_fix, iters)
        Tb, db = backend_gn_odom(ref_backend, g, od, dev, num_fix, iters)
        assert rel_err(Ta, Tb) < GN_TOL, (num_fix, iters)
        if iters == 1:
            assert rel_err(da, db) < GN_TOL
    capfd.readouterr()


def test_refine_patch_variants_agree_bitwise(be, dev):
    """refine_matches runs 16x16 patches (two CTAs per SM) for F <= 24 and keeps the 32x32 kernel
    for F = 32; both must give the same bits.  frodo_set_refine_patch forces the variant."""
    from frodo_slam_b200 import _lib
    from frodo_slam_b200 import synth
    mp = synth.MATCH_PARAMS
    B, H, W = 2, 384, 512
    d = synth.matching_pairs(B, H, W, seed=4, device=dev)
    rays, pts, p0 = synth.prep_for_iter_proj(d["X11"], d["X21"])
    p1 = be.iter_proj(rays, pts, p0, mp["max_iter"], mp["lambda_init"], mp["convergence_thresh"])[0].long()
    g = torch.Generator(device=dev).manual_seed(1)
    p1 = p1 + torch.randint(-6, 7, p1.shape, device=dev, generator=g)  # off-box candidates too
    D11 = d["D11"].half().contiguous()
    D21 = d["D21"].half().reshape(B, H * W, -1).contiguous()
    out = {}
    try:
        for v in ("16", "32"):
            assert _lib.load().frodo_set_refine_patch(int(v)) == int(v)
            pn, sc = be.refine_matches(D11, D21, p1, mp["radius"], mp["dilation_max"])
            torch.cuda.synchronize()
            out[v] = (pn.clone(), sc.clone())
    finally:
        _lib.load().frodo_set_refine_patch(0)
    assert torch.equal(out["16"][0], out["32"][0])
    assert torch.equal(out["16"][1].view(torch.int16), out["32"][1].view(torch.int16))
