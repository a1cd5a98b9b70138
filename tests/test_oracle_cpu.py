"""Self-consistency of the CPU oracle (independent of the golden vectors)."""
import ctypes

import numpy as np
import torch

from helpers import GP, MP, graph_inputs, matching_inputs, oracle_gn, gn_args


def test_half_conversion_matches_numpy():
    from oracle import oracle as o
    rng = np.random.default_rng(0)
    x = (rng.standard_normal(200000) * 10.0 ** rng.integers(-9, 5, 200000)).astype(np.float32)
    x = np.concatenate([x, np.array([0, -0.0, 65504, 65519.9, 65520, 1e-8, 2.0 ** -25, 2.0 ** -24,
                                     np.inf, -np.inf], np.float32)])
    out = np.zeros(x.size, np.uint16)
    o.lib().oracle_f2h(o._p(x), o._p(out), ctypes.c_int64(x.size))
    with np.errstate(over="ignore"):
        assert np.array_equal(out, x.astype(np.float16).view(np.uint16))
    h = np.arange(65536, dtype=np.uint16)
    f = np.zeros(65536, np.float32)
    o.lib().oracle_h2f(o._p(h), o._p(f), ctypes.c_int64(65536))
    r = h.view(np.float16).astype(np.float32)
    assert np.all((f == r) | (np.isnan(f) & np.isnan(r)))


def test_refine_oracle_vs_numpy_float16_bruteforce():
    from oracle import oracle as o
    rng = np.random.default_rng(1)
    H, W, F = 12, 14, 24
    D11 = rng.standard_normal((1, H, W, F)).astype(np.float16)
    D21 = rng.standard_normal((1, H * W, F)).astype(np.float16)
    p1 = np.stack([rng.integers(-3, W + 3, H * W), rng.integers(-3, H + 3, H * W)], -1)[None]
    pn, sc = o.refine_matches(D11, D21, p1.astype(np.int64), 2, 3)
    for n in range(0, H * W, 7):
        u0, v0 = p1[0, n]
        best, bu, bv = np.float16(0), u0, v0
        for d in (3, 2, 1):
            for i in range(0, 4 * d + 1, d):
                for j in range(0, 4 * d + 1, d):
                    u, v = u0 - 2 * d + i, v0 - 2 * d + j
                    if 0 <= u < W and 0 <= v < H:
                        s = np.float16(0)
                        for k in range(F):
                            s = np.float16(s + np.float16(D21[0, n, k] * D11[0, v, u, k]))
                        if s > best:
                            best, bu, bv = s, u, v
            u0, v0 = bu, bv
        assert (pn[0, n, 0], pn[0, n, 1]) == (bu, bv)
        assert sc[0, n] == best


def test_iter_proj_oracle_converges_to_analytic_match():
    from oracle import oracle as o
    d, rays, pts, p0 = matching_inputs(1, 48, 64, seed=2)
    p, conv = o.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), MP["max_iter"], MP["lambda_init"],
                          MP["convergence_thresh"])
    pt = d["p_true"].numpy()
    inside = (pt[..., 0] > 2) & (pt[..., 0] < 61) & (pt[..., 1] > 2) & (pt[..., 1] < 45)
    err = np.abs(p - pt).max(-1)
    assert conv[inside].mean() > 0.99
    assert np.median(err[inside]) < 0.01


def test_chol_solve_matches_numpy_and_reports_failure():
    from oracle import oracle as o
    rng = np.random.default_rng(3)
    A = rng.standard_normal((40, 40))
    H = A @ A.T + 40 * np.eye(40)
    g = rng.standard_normal(40)
    x, fail = o.chol_solve(H, g)
    assert not fail and np.allclose(x, np.linalg.solve(H, g), rtol=1e-10, atol=1e-12)
    x, fail = o.chol_solve(-H, g)
    assert fail and np.all(x == 0)


def test_edge_blocks_structure():
    """Ji = -Jj (gn_kernels.cu:769): H_ii = H_jj = -H_ij = -H_ji^T, g_i = -g_j."""
    from oracle import oracle as o
    g = graph_inputs("window8", 24, 32)
    ie, je, nu = o.edge_index(g["ii"].numpy(), g["jj"].numpy())
    assert nu == 8
    for mode in ("rays", "points", "calib"):
        kw, _ = gn_args(mode, g)
        okw = {k: v for k, v in kw.items() if k not in ("sigma_a", "sigma_b")}
        Hs, gs = o.edge_blocks(mode, g["Twc"].numpy(), g["Xs"].numpy(), g["Cs"].numpy(), ie, je,
                               g["idx_ii2jj"].numpy(), g["valid_match"].numpy(), g["Q"].numpy(),
                               kw["sigma_a"], kw["sigma_b"], GP["C_conf"], GP["Q_conf"], **okw)
        tol = dict(rtol=1e-5, atol=1e-6 * np.abs(Hs[3]).max())  # fp32 product order differs
        assert np.allclose(Hs[0], Hs[3], **tol) and np.allclose(Hs[1], -Hs[3], **tol)
        assert np.allclose(Hs[2], -Hs[3], **tol) and np.allclose(gs[0], -gs[1], rtol=1e-12)
        assert np.abs(Hs[3]).max() > 0


def test_edge_index_arbitrary_ids():
    from oracle import oracle as o
    ii = np.array([50, 7, 7, 1000, 50], np.int64)
    jj = np.array([7, 50, 1000, 7, 3], np.int64)
    ie, je, nu = o.edge_index(ii, jj)
    uniq = np.unique(np.concatenate([ii, jj]))
    assert nu == len(uniq)
    assert np.array_equal(ie, np.searchsorted(uniq, ii)) and np.array_equal(je, np.searchsorted(uniq, jj))


def test_gn_oracle_reduces_cost_and_honours_delta_thresh():
    g = graph_inputs("window8", 24, 32)
    r = oracle_gn("points", g, 1, 10, delta_thresh=1e-3)
    assert 1 <= r["iters"] < 10 and r["fails"] == 0
    e0 = np.abs(g["Twc"].numpy() - g["Twc_true"].numpy()).max()
    e1 = np.abs(r["Twc"] - g["Twc_true"].numpy()).max()
    assert e1 < e0
    assert np.array_equal(r["Twc"][0], g["Twc"].numpy()[0])  # fixed pose untouched


def test_odometry_blocks_structure():
    """odom restatement: zero residual at a consistent measurement, symmetric PSD 14x14 block,
    J_i = -J_j except for the scale column (gn_kernels.cu:1236-1241)."""
    import pytest
    from frodo_slam_b200 import synth
    from helpers import OP
    from oracle import oracle
    g = graph_inputs("window8", 16, 16)
    od = synth.odometry_edges(g, noise_t=0.0, noise_r=0.0)
    Hs, gs = oracle.odom_blocks(g["Twc_true"].numpy(), od["Tij"].numpy(), od["odom_ii"].numpy(),
                                od["odom_jj"].numpy(), OP["sigma_odom_t"], OP["sigma_odom_r"])
    info = 1.0 / OP["sigma_odom_t"] ** 2
    assert np.abs(gs).max() < 2e-3 * info  # residual ~ fp32 rounding of the relative pose
    for e in range(Hs.shape[1]):
        full = np.block([[Hs[0, e], Hs[1, e]], [Hs[2, e], Hs[3, e]]]).astype(np.float64)
        assert np.allclose(full, full.T, rtol=1e-5, atol=1e-5 * info)
        assert np.linalg.eigvalsh(0.5 * (full + full.T)).min() > -1e-4 * info
        assert np.allclose(Hs[0, e][:6, :6], Hs[3, e][:6, :6], rtol=1e-5, atol=1e-6 * info)
        assert np.all(Hs[0, e][6, :] == 0) and np.all(Hs[0, e][:, 6] == 0)
    H = np.zeros((49, 49))
    gv = np.zeros(49)
    oracle.scale_prior(H, gv, g["Twc"].numpy(), od["s_bar"].numpy(), 1, 0.1)
    assert np.count_nonzero(H) == 6 and H[6, 6] == pytest.approx(100.0)  # keyframe 2 skipped


def test_match_quality_oracle_equals_torch_expressions():
    """tracker.py:194-206 / global_opt.py:98-112 evaluated by torch (CPU) vs the numpy restatement."""
    import torch
    from oracle import oracle
    g = torch.Generator().manual_seed(3)
    B, N, Ns = 3, 501, 377
    Qs, Qd = torch.rand(B, Ns, 1, generator=g) * 3, torch.rand(B, N, 1, generator=g) * 3
    Cs, Cd = torch.rand(B, Ns, 1, generator=g) - 0.3, torch.rand(B, N, 1, generator=g) - 0.3
    idx = torch.randint(0, Ns, (B, N), generator=g)
    vm = torch.rand(B, N, 1, generator=g) < 0.7
    rows = torch.arange(B)[:, None]
    Q = torch.sqrt(Qs[rows, idx] * Qd)
    vq = vm & (Q > 1.5)
    vo = vq & (Cs[rows, idx] > 0.1) & (Cd > 0.1)
    o = oracle.match_quality(Qs.numpy(), Qd.numpy(), idx.numpy(), vm.numpy(), 1.5, Cs.numpy(),
                             Cd.numpy(), 0.1)
    # torch's CPU sqrt is a vectorised approximation (1 ulp off in ~0.6 % of the values); numpy's,
    # like CUDA's sqrtf the reference runs on, is correctly rounded: compare to 1 ulp, and the
    # masks away from the threshold
    Qn = Q.numpy().reshape(B, N)
    assert np.abs(o[0] - Qn).max() <= 2.4e-7 * 3
    clear = np.abs(Qn - 1.5) > 1e-5
    assert np.array_equal(o[1][clear], vo.numpy().reshape(B, N)[clear])
    assert np.array_equal(o[2][clear], vq.numpy().reshape(B, N)[clear])
    assert o[3].tolist() == [[int(o[1][b].sum()), int(o[2][b].sum())] for b in range(B)]
