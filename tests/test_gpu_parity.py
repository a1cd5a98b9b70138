"""GPU parity tests: the CUDA path (through the C ABI via mast3r_slam_backends) against the
CPU oracle, and against the unmodified reference backend (oracle/_ref) when it is present.

Bars (BASELINE.json north_star): match indices / masks / scores bit-exact; iter_proj within
1e-4 px (we require bit-exact); H, g, dx and poses within 1e-4 relative (norm-wise)."""
import numpy as np
import pytest
import torch

from helpers import (GP, MP, backend_gn, expand_edge_blocks, graph_inputs, matching_inputs,
                     oracle_gn, rel_err)

pytestmark = pytest.mark.gpu

GN_TOL = 1e-4  # relative, norm-wise (north_star)


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def be():
    import mast3r_slam_backends
    return mast3r_slam_backends


# ------------------------------------------------------------------ iter_proj

@pytest.mark.parametrize("B,H,W,init", [(1, 48, 64, "identity"), (3, 37, 53, "identity"),
                                        (2, 48, 64, "random"), (2, 48, 64, "outside"),
                                        (1, 224, 224, "identity")])
def test_iter_proj_bit_exact_vs_oracle(be, dev, B, H, W, init):
    from oracle import oracle
    _, rays, pts, p0 = matching_inputs(B, H, W, seed=3, init=init)
    p, c = be.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), MP["max_iter"], MP["lambda_init"],
                        MP["convergence_thresh"])
    pr, cr = oracle.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), MP["max_iter"],
                              MP["lambda_init"], MP["convergence_thresh"])
    assert c.dtype == torch.bool and p.dtype == torch.float32
    diff = np.abs(p.cpu().numpy() - pr).max()
    assert diff <= 1e-4, f"max |dp| = {diff} px"  # north_star tolerance
    assert np.array_equal(p.cpu().numpy(), pr), "not bit-exact"
    assert np.array_equal(c.cpu().numpy(), cr)


@pytest.mark.parametrize("max_iter", [0, 1, 4])
def test_iter_proj_iteration_counts(be, dev, max_iter):
    from oracle import oracle
    _, rays, pts, p0 = matching_inputs(1, 40, 56, seed=5)
    p, c = be.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), max_iter, 1e-8, 1e-6)
    pr, cr = oracle.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), max_iter, 1e-8, 1e-6)
    assert np.array_equal(p.cpu().numpy(), pr) and np.array_equal(c.cpu().numpy(), cr)


def test_iter_proj_vs_reference(be, dev, ref_backend):
    if ref_backend is None:
        pytest.skip("oracle/_ref not built")
    _, rays, pts, p0 = matching_inputs(2, 64, 80, seed=7)  # N % 16 == 0 (reference has no guard)
    a = be.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), 10, 1e-8, 1e-6)
    b = ref_backend.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), 10, 1e-8, 1e-6)
    torch.cuda.synchronize()
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


# ------------------------------------------------------------- refine_matches

def _refine_inputs(B, H, W, seed, dtype=torch.float16, F=24):
    d = synth_pairs(B, H, W, seed, F)
    g = torch.Generator().manual_seed(seed)
    # centres: true match + up to 4 px error, some far outside the image
    p = d["p_true"].round().long() + torch.randint(-4, 5, (B, H * W, 2), generator=g)
    p[:, ::97] += 40
    p[:, 5::101] -= 60
    D11 = d["D11"].to(dtype).contiguous()
    D21 = d["D21"].to(dtype).reshape(B, H * W, -1).contiguous()
    return D11, D21, p.contiguous()


def synth_pairs(B, H, W, seed, F=24):
    from frodo_slam_b200 import synth
    return synth.matching_pairs(B, H, W, seed=seed, F=F)


@pytest.mark.parametrize("B,H,W,F,dtype", [(1, 48, 64, 24, torch.float16),
                                           (2, 37, 53, 24, torch.float16),
                                           (1, 40, 40, 16, torch.float16),
                                           (1, 40, 40, 20, torch.float16),
                                           (1, 32, 48, 24, torch.float32)])
def test_refine_matches_bit_exact_vs_oracle(be, dev, B, H, W, F, dtype):
    from oracle import oracle
    D11, D21, p1 = _refine_inputs(B, H, W, 11, dtype, F)
    pn, sc = be.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), MP["radius"],
                               MP["dilation_max"])
    pr, sr = oracle.refine_matches(D11.numpy(), D21.numpy(), p1.numpy(), MP["radius"],
                                   MP["dilation_max"])
    assert np.array_equal(pn.cpu().numpy(), pr)
    w = np.uint16 if dtype == torch.float16 else np.uint32
    assert np.array_equal(sc.cpu().numpy().view(w), sr.view(w))


def test_refine_matches_float64(be, dev):
    """The reference dispatches double as well (matching_kernels.cu:112): same indices as a
    float64 brute force in torch (fma order as the kernel: k ascending)."""
    D11, D21, p1 = _refine_inputs(1, 24, 32, 19, torch.float64)
    pn, sc = be.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), 2, 2)
    assert sc.dtype == torch.float64 and pn.dtype == torch.int64
    H, W = 24, 32
    q = D21[0]
    best = torch.full((H * W,), 2.2250738585072014e-308, dtype=torch.float64)
    u0, v0 = p1[0, :, 0].clone(), p1[0, :, 1].clone()
    bu, bv = u0.clone(), v0.clone()
    img = D11[0].reshape(H * W, -1)
    for d in (2, 1):
        rd = 2 * d
        for i in range(0, 2 * rd + 1, d):
            for j in range(0, 2 * rd + 1, d):
                u, v = u0 - rd + i, v0 - rd + j
                inside = (v >= 0) & (v < H) & (u >= 0) & (u < W)
                c = img[v.clamp(0, H - 1) * W + u.clamp(0, W - 1)]
                s = (q * c).sum(1)
                better = inside & (s > best + 1e-12)
                best = torch.where(better, s, best)
                bu, bv = torch.where(better, u, bu), torch.where(better, v, bv)
        u0, v0 = bu.clone(), bv.clone()
    same = (pn[0, :, 0].cpu() == bu) & (pn[0, :, 1].cpu() == bv)
    assert same.float().mean() > 0.995  # ties at the 1e-12 level may resolve differently
    assert torch.allclose(sc[0].cpu()[same], best[same], rtol=1e-10, atol=1e-300)


@pytest.mark.parametrize("radius,dil", [(1, 1), (2, 3), (3, 5), (0, 2)])
def test_refine_matches_search_shapes(be, dev, radius, dil):
    from oracle import oracle
    D11, D21, p1 = _refine_inputs(1, 32, 40, 13)
    pn, sc = be.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), radius, dil)
    pr, sr = oracle.refine_matches(D11.numpy(), D21.numpy(), p1.numpy(), radius, dil)
    assert np.array_equal(pn.cpu().numpy(), pr)
    assert np.array_equal(sc.cpu().numpy().view(np.uint16), sr.view(np.uint16))


def test_refine_matches_vs_reference(be, dev, ref_backend):
    if ref_backend is None:
        pytest.skip("oracle/_ref not built")
    D11, D21, p1 = _refine_inputs(2, 64, 80, 17)
    a = be.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), 3, 5)
    b = ref_backend.refine_matches(D11.to(dev), D21.to(dev), p1.to(dev), 3, 5)
    torch.cuda.synchronize()
    assert torch.equal(a[0], b[0])
    assert torch.equal(a[1].view(torch.int16), b[1].view(torch.int16))


def test_match_end_to_end_vs_reference(be, dev, ref_backend):
    """matching.py:52-92 composed from the two ops: idx_1_to_2 and valid bit-equal."""
    if ref_backend is None:
        pytest.skip("oracle/_ref not built")
    d, rays, pts, p0 = matching_inputs(2, 64, 80, seed=19)
    outs = []
    for m in (be, ref_backend):
        p1, conv = m.iter_proj(rays.to(dev), pts.to(dev), p0.to(dev), 10, 1e-8, 1e-6)
        p1 = p1.long()
        pn, sc = m.refine_matches(d["D11"].half().to(dev),
                                  d["D21"].half().reshape(2, 64 * 80, -1).to(dev), p1, 3, 5)
        valid = conv & (sc > MP["ft_matching_threshold"])
        outs.append((pn[..., 0] + 80 * pn[..., 1], valid))
    torch.cuda.synchronize()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


# --------------------------------------------------------------- Gauss-Newton

@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
def test_gn_system_vs_oracle(dev, mode):
    """Per-edge blocks and the assembled H, g of one accumulation pass."""
    from frodo_slam_b200 import backends
    from oracle import oracle
    from helpers import gn_args
    g = graph_inputs("window8", 48, 64, id_base=7, id_stride=3)
    kw, _ = gn_args(mode, g)
    prm = dict(rays=dict(sigma_ray=kw["sigma_a"], sigma_dist=kw["sigma_b"]),
               points=dict(sigma_point=kw["sigma_a"]),
               calib=dict(sigma_pixel=kw["sigma_a"], sigma_depth=kw["sigma_b"], height=g["H"],
                          width=g["W"], pixel_border=GP["pixel_border"], z_eps=GP["depth_eps"]))
    p = backends.make_params(mode, C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"], **prm[mode])
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    H, gv, eb, ii_row, jj_row = backends.gn_system(
        p, t["Twc"], t["Xs"], t["Cs"], t["K"] if mode == "calib" else None, t["ii"], t["jj"],
        t["idx_ii2jj"], t["valid_match"], t["Q"], 1)
    torch.cuda.synchronize()
    ie, je, nu = oracle.edge_index(g["ii"].numpy(), g["jj"].numpy())
    assert np.array_equal(ii_row.cpu().numpy(), ie) and np.array_equal(jj_row.cpu().numpy(), je)
    okw = {k: v for k, v in kw.items() if k not in ("sigma_a", "sigma_b")}
    Hs, gs = oracle.edge_blocks(mode, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie, je,
                                g["idx_ii2jj"].numpy(), g["valid_match"].numpy(), g["Q"].numpy(),
                                kw["sigma_a"], kw["sigma_b"], GP["C_conf"], GP["Q_conf"], **okw)
    Hs_o, gs_o = expand_edge_blocks(eb.cpu().numpy())
    for e in range(Hs.shape[1]):
        assert rel_err(Hs_o[:, e], Hs[:, e]) < GN_TOL, f"edge {e} H blocks"
        assert rel_err(gs_o[:, e], gs[:, e]) < GN_TOL, f"edge {e} g blocks"
    Hr, gr = oracle.assemble(Hs, gs, ie, je, g["Xs"].shape[0], 1, round_f32=False)
    assert rel_err(H.cpu().numpy(), Hr) < GN_TOL
    assert rel_err(gv.cpu().numpy(), gr) < GN_TOL
    assert np.allclose(H.cpu().numpy(), H.cpu().numpy().T, rtol=0, atol=0)


@pytest.mark.parametrize("mode,cfg,num_fix,iters", [("rays", "pair", 1, 1), ("rays", "window8", 1, 4),
                                                    ("rays", "window8", 2, 3),
                                                    ("points", "window8", 1, 4),
                                                    ("calib", "window8", 1, 4),
                                                    ("rays", "calib32", 1, 2)])
def test_gn_solve_vs_oracle(be, dev, mode, cfg, num_fix, iters):
    g = graph_inputs(cfg, 48, 64)
    Twc, dx = backend_gn(be, mode, g, dev, num_fix, iters)
    ref = oracle_gn(mode, g, num_fix, iters)
    assert ref["fails"] == 0
    assert rel_err(dx, ref["dx"]) < 5e-3  # dx of a late iteration amplifies by cond(H); poses are the gate
    assert rel_err(Twc, ref["Twc"]) < GN_TOL


def test_gn_first_iteration_dx(be, dev):
    g = graph_inputs("window8", 48, 64)
    for mode in ("rays", "points", "calib"):
        _, dx = backend_gn(be, mode, g, dev, 1, 1)
        ref = oracle_gn(mode, g, 1, 1)
        assert rel_err(dx, ref["dx"]) < GN_TOL, mode


def test_gn_convergence_break(be, dev):
    """delta_thresh stops the iteration on the device exactly where the oracle stops."""
    g = graph_inputs("window8", 48, 64)
    ref = oracle_gn("points", g, 1, 10, delta_thresh=1e-3)
    assert 1 < ref["iters"] < 10
    Twc, dx = backend_gn(be, "points", g, dev, 1, 10, delta_thresh=1e-3)
    assert rel_err(Twc, ref["Twc"]) < GN_TOL and rel_err(dx, ref["dx"]) < 1e-2


def test_gn_not_positive_definite_returns_zero(be, dev):
    """All matches invalid -> H = 0 -> Cholesky fails -> dx = 0, poses untouched (gn_kernels.cu:193-197)."""
    g = graph_inputs("window8", 32, 40)
    g["valid_match"] = torch.zeros_like(g["valid_match"])
    Twc, dx = backend_gn(be, "rays", g, dev, 1, 3)
    assert np.all(dx == 0) and np.array_equal(Twc, g["Twc"].numpy())


def test_gn_vs_reference(be, dev, ref_backend, capfd):
    if ref_backend is None:
        pytest.skip("oracle/_ref not built")
    g = graph_inputs("window8", 48, 64)
    for mode in ("rays", "points", "calib"):
        for iters in (1, 3):
            Ta, da = backend_gn(be, mode, g, dev, 1, iters)
            Tb, db = backend_gn(ref_backend, mode, g, dev, 1, iters)
            assert rel_err(Ta, Tb) < GN_TOL, (mode, iters)
            if iters == 1:
                assert rel_err(da, db) < GN_TOL, (mode, iters)
    capfd.readouterr()  # the reference prints its solver log to stdout
