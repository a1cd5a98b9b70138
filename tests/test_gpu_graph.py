"""GPU: the device-resident keyframe / edge arena (replaces the per-solve
stacking of global_opt.py:186-224) against the drop-in operators on the stacked tensors.

The arena runs the same kernels with the same edge order, so poses and dx must be
BIT-IDENTICAL to `gauss_newton_*` on `stacked_inputs()`; against the oracle (which sums in
the reference's order) the usual 1e-4 bar applies."""
import numpy as np
import pytest
import torch

from oracle import oracle
from frodo_slam_b200 import synth

pytestmark = pytest.mark.gpu
GP = synth.GN_PARAMS


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _build(dev, cfg, H, W, id_base, Kcap, Ecap, batches=2, seed=1):
    """Fill an arena from a synthetic graph: keyframe k gets id id_base + k."""
    from frodo_slam_b200.graph import KeyframeGraphArena
    K, und = synth.graph_edges(cfg)
    g = synth.keyframe_graph(K, H, W, und, seed=seed, device="cpu")
    g = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    N = H * W
    A = KeyframeGraphArena(N, Kcap, Ecap, device=dev)
    for k in range(K):
        A.put_keyframe(id_base + k, g["Xs"][k], g["Cs"][k], g["Twc"][k])
    nu = len(und)
    idx, valid, Q = g["idx_ii2jj"], g["valid_match"], g["Q"]
    for b in range(batches):
        sel = list(range(b, nu, batches))
        A.add_edges([id_base + und[e][0] for e in sel], [id_base + und[e][1] for e in sel],
                    idx[sel], idx[[nu + e for e in sel]], valid[sel], valid[[nu + e for e in sel]],
                    Q[sel], Q[[nu + e for e in sel]])
    return A, g, und


def _dropin(mode, st, num_fix, iters, g):
    import mast3r_slam_backends as be
    uniq, Twc, Xs, Cs, ii, jj, idx, valid, Q = st
    Twc = Twc.clone()
    if mode == "rays":
        dx = be.gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, GP["sigma_ray"],
                                  GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], num_fix, iters, 0.0)[0]
    elif mode == "points":
        dx = be.gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, GP["sigma_point"],
                                    GP["C_conf"], GP["Q_conf"], iters, 0.0)[0]
    else:
        dx = be.gauss_newton_calib(Twc, Xs, Cs, g["K"], ii, jj, idx, valid, Q, g["H"], g["W"],
                                   GP["pixel_border"], GP["depth_eps"], GP["sigma_pixel"],
                                   GP["sigma_depth"], GP["C_conf"], GP["Q_conf"], iters, 0.0)[0]
    return Twc, dx


def _solve(A, mode, num_fix, iters, g):
    if mode == "rays":
        return A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], num_fix,
                            iters, 0.0)
    if mode == "points":
        return A.solve("points", 1, iters, 0.0, sigma_point=GP["sigma_point"],
                       C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"])
    return A.solve("calib", 1, iters, 0.0, K=g["K"], sigma_pixel=GP["sigma_pixel"],
                   sigma_depth=GP["sigma_depth"], C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"],
                   height=g["H"], width=g["W"], pixel_border=GP["pixel_border"],
                   z_eps=GP["depth_eps"])


@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
@pytest.mark.parametrize("H,W", [(48, 64), (37, 53)])
def test_arena_solve_equals_dropin_bitwise(dev, mode, H, W):
    # ids 10..17 in a ring of 12 slots: slots wrap (10, 11, 0, 1, ...), so arena rows != ranks
    A, g, _ = _build(dev, "window8", H, W, id_base=10, Kcap=12, Ecap=16)
    st = A.stacked_inputs()
    Twc_ref, dx_ref = _dropin(mode, st, 1, 3, g)
    dx, uniq = _solve(A, mode, 1, 3, g)
    assert uniq == list(range(10, 18))
    assert torch.equal(dx, dx_ref)
    assert torch.equal(A.poses(uniq), Twc_ref)


def test_arena_vs_oracle_reference_order(dev):
    """Against the CPU oracle on the reference's own stacking (all forward edges, then all
    backward edges, global_opt.py:205-210): 1e-4 on the poses."""
    A, g, und = _build(dev, "window8", 48, 64, id_base=0, Kcap=8, Ecap=12)
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 3, 0.0)
    c = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in g.items()}
    ref = oracle.gauss_newton("rays", c["Twc"], c["Xs"], c["Cs"], c["ii"], c["jj"], c["idx_ii2jj"],
                              c["valid_match"], c["Q"], GP["sigma_ray"], GP["sigma_dist"],
                              GP["C_conf"], GP["Q_conf"], 1, 3, 0.0)
    assert np.abs(A.poses(uniq).cpu().numpy() - ref["Twc"]).max() < 1e-4


def test_num_fix_2_and_eviction(dev):
    A, g, und = _build(dev, "window8", 48, 64, id_base=20, Kcap=8, Ecap=12)
    assert A.num_edges() == 12
    # slide the window: keyframes 20, 21 leave (and every edge touching them)
    assert A.evict_older_than(22)
    ii, jj = A.edge_list()
    assert all(i >= 22 and j >= 22 for i, j in zip(ii, jj))
    assert sorted(zip(ii, jj)) == sorted((20 + i, 20 + j) for i, j in und if i >= 2 and j >= 2)
    assert A.keyframe_ids() == list(range(22, 28))
    st = A.stacked_inputs()
    Twc_ref, dx_ref = _dropin("rays", st, 2, 3, g)
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 2, 3, 0.0)
    assert uniq == list(range(22, 28)) and dx.shape == (4, 7)
    assert torch.equal(dx, dx_ref) and torch.equal(A.poses(uniq), Twc_ref)
    # the freed ring slots take new keyframes (ids 28, 29 -> slots 4, 5) and a new edge
    A.put_keyframe(28, g["Xs"][0], g["Cs"][0], g["Twc"][0])
    A.put_keyframe(29, g["Xs"][1], g["Cs"][1], g["Twc"][1])
    nu = len(und)
    A.add_edges([28], [29], g["idx_ii2jj"][0:1], g["idx_ii2jj"][nu:nu + 1], g["valid_match"][0:1],
                g["valid_match"][nu:nu + 1], g["Q"][0:1], g["Q"][nu:nu + 1])
    st = A.stacked_inputs()
    assert st[0] == list(range(22, 30))
    Twc_ref, dx_ref = _dropin("rays", st, 1, 2, g)
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 2, 0.0)
    assert torch.equal(dx, dx_ref) and torch.equal(A.poses(uniq), Twc_ref)


def test_keyframe_without_edges_is_not_in_the_system(dev):
    """unique_kf_idx comes from the edges only (global_opt.py:180-184)."""
    A, g, _ = _build(dev, "pair", 32, 48, id_base=3, Kcap=6, Ecap=4, batches=1)
    A.put_keyframe(5, g["Xs"][0], g["Cs"][0], g["Twc"][0])  # no edge touches it
    before = A.poses([5]).clone()
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 2, 0.0)
    assert uniq == [3, 4] and dx.shape == (1, 7)
    assert torch.equal(A.poses([5]), before)
    Twc_ref, dx_ref = _dropin("rays", A.stacked_inputs(), 1, 2, g)
    # the solve above already moved the poses; redo from a fresh arena for the comparison
    B, g2, _ = _build(dev, "pair", 32, 48, id_base=3, Kcap=6, Ecap=4, batches=1)
    Twc_ref, dx_ref = _dropin("rays", B.stacked_inputs(), 1, 2, g2)
    assert torch.equal(dx, dx_ref)


def test_pointmap_update_repacks_the_slot(dev):
    """tracker.py:318-320: the last keyframe's pointmap changes after every tracked frame."""
    A, g, _ = _build(dev, "window8", 48, 64, id_base=0, Kcap=8, Ecap=12)
    X_new = (g["Xs"][7] * 1.01).contiguous()
    A.put_keyframe(7, X_new, g["Cs"][7])
    st = A.stacked_inputs()
    assert torch.equal(st[2][7], X_new)
    Twc_ref, dx_ref = _dropin("rays", st, 1, 2, g)
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 2, 0.0)
    assert torch.equal(dx, dx_ref) and torch.equal(A.poses(uniq), Twc_ref)


def test_capacity_and_argument_errors(dev):
    from frodo_slam_b200.graph import KeyframeGraphArena
    A, g, _ = _build(dev, "pair", 16, 24, id_base=0, Kcap=2, Ecap=1, batches=1)
    with pytest.raises(RuntimeError):  # slot 0 still holds keyframe 0
        A.put_keyframe(2, g["Xs"][0], g["Cs"][0], g["Twc"][0])
    with pytest.raises(RuntimeError):  # edge capacity
        A.add_edges([0], [1], g["idx_ii2jj"][0:1], g["idx_ii2jj"][1:2], g["valid_match"][0:1],
                    g["valid_match"][1:2], g["Q"][0:1], g["Q"][1:2])
    with pytest.raises(KeyError):
        A.poses([9])
    with pytest.raises(RuntimeError):  # CPU tensor: no CPU path
        A.put_keyframe(1, g["Xs"][1].cpu(), g["Cs"][1], None)
    with pytest.raises(RuntimeError):
        KeyframeGraphArena(16, 2, 1, device="cpu")
    assert not A.evict_older_than(5) and A.num_edges() == 0 and A.keyframe_ids() == []
    with pytest.raises(RuntimeError):
        A.solve_rays(0.003, 10.0, 0.0, 1.5, 1, 1, 0.0)


def test_remove_keyframe_and_reinsert(dev):
    """FactorGraph.remove_factors_i: the newest keyframe is replaced together with its edges."""
    A, g, und = _build(dev, "window8", 48, 64, id_base=0, Kcap=8, Ecap=12)
    nu = len(und)
    sel = [e for e in range(nu) if 7 in und[e]]
    A.remove_keyframe(7)
    assert A.num_edges() == nu - len(sel) and 7 not in A.keyframe_ids()
    A.put_keyframe(7, g["Xs"][7], g["Cs"][7], g["Twc"][7])
    A.add_edges([und[e][0] for e in sel], [und[e][1] for e in sel], g["idx_ii2jj"][sel],
                g["idx_ii2jj"][[nu + e for e in sel]], g["valid_match"][sel],
                g["valid_match"][[nu + e for e in sel]], g["Q"][sel], g["Q"][[nu + e for e in sel]])
    assert sorted(zip(*A.edge_list())) == sorted(und)
    st = A.stacked_inputs()
    Twc_ref, dx_ref = _dropin("rays", st, 1, 3, g)
    dx, uniq = A.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 3, 0.0)
    assert torch.equal(dx, dx_ref) and torch.equal(A.poses(uniq), Twc_ref)
    # and the result agrees with the untouched graph to rounding (same terms, other order)
    B, _, _ = _build(dev, "window8", 48, 64, id_base=0, Kcap=8, Ecap=12)
    B.solve_rays(GP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], 1, 3, 0.0)
    assert (A.poses(uniq) - B.poses(uniq)).abs().max().item() < 1e-5


# ---------------------------------------------- gauss_newton_rays_odom through the arena / dist

OP = synth.ODOM_PARAMS


@pytest.mark.parametrize("id_base,Kcap", [(0, 8), (10, 12)])
@pytest.mark.parametrize("num_fix", [1, 2])
def test_arena_rays_odom_equals_dropin_bitwise(dev, id_base, Kcap, num_fix):
    import mast3r_slam_backends as be
    A, g, und = _build(dev, "window8", 48, 64, id_base=id_base, Kcap=Kcap, Ecap=12)
    od = synth.odometry_edges(g, device=dev)
    uniq, Twc, Xs, Cs, ii, jj, idx, valid, Q = A.stacked_inputs()
    # the drop-in uses ids as rows of the stacked Twc (gn_kernels.cu:1121-1122): hand it ranks
    base = torch.tensor(uniq, device=dev)
    ii_r, jj_r = torch.searchsorted(base, ii), torch.searchsorted(base, jj)
    Twc_ref = Twc.clone()
    dx_ref = be.gauss_newton_rays_odom(
        Twc_ref, Xs, Cs, ii_r, jj_r, idx, valid, Q, od["odom_ii"], od["odom_jj"], od["Tij"],
        od["s_bar"], OP["sigma_odom_t"], OP["sigma_odom_r"], OP["sigma_ray"], GP["sigma_dist"],
        OP["sigma_scale_prior"], GP["C_conf"], GP["Q_conf"], num_fix, 3, 0.0)[0]
    dx, uniq2 = A.solve_rays_odom(od["odom_ii"] + id_base, od["odom_jj"] + id_base, od["Tij"],
                                  od["s_bar"], OP["sigma_odom_t"], OP["sigma_odom_r"],
                                  OP["sigma_ray"], GP["sigma_dist"], OP["sigma_scale_prior"],
                                  GP["C_conf"], GP["Q_conf"], num_fix, 3, 0.0)
    assert uniq2 == uniq
    assert torch.equal(dx, dx_ref) and torch.equal(A.poses(uniq), Twc_ref)
    # and the odometry terms do change the result
    B, _, _ = _build(dev, "window8", 48, 64, id_base=id_base, Kcap=Kcap, Ecap=12)
    B.solve_rays(OP["sigma_ray"], GP["sigma_dist"], GP["C_conf"], GP["Q_conf"], num_fix, 3, 0.0)
    assert (B.poses(uniq) - A.poses(uniq)).abs().max().item() > 1e-6


def test_sharded_driver_with_odometry_equals_single_call_bitwise(dev):
    """dist.ShardedGaussNewton.set_odometry (terms added after the all-reduce) at world 1."""
    import mast3r_slam_backends as be
    from frodo_slam_b200 import backends, dist as fdist
    K, und = synth.graph_edges("window8")
    g = synth.keyframe_graph(K, 48, 64, und, seed=3, device="cpu")
    g = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    od = synth.odometry_edges(g, device=dev)
    E, N = g["ii"].shape[0], 48 * 64
    Ta = g["Twc"].clone()
    da = be.gauss_newton_rays_odom(
        Ta, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"],
        od["odom_ii"], od["odom_jj"], od["Tij"], od["s_bar"], OP["sigma_odom_t"],
        OP["sigma_odom_r"], OP["sigma_ray"], GP["sigma_dist"], OP["sigma_scale_prior"],
        GP["C_conf"], GP["Q_conf"], 1, 3, 0.0)[0]
    prm = backends.make_params("rays", sigma_ray=OP["sigma_ray"], sigma_dist=GP["sigma_dist"],
                               C_thresh=GP["C_conf"], Q_thresh=GP["Q_conf"])
    Tb = g["Twc"].clone()
    sg = fdist.ShardedGaussNewton(prm, Tb, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                  g["idx_ii2jj"].view(E, N), g["valid_match"].view(E, N),
                                  g["Q"].view(E, N), 0, E, 1)
    sg.set_odometry(od["odom_ii"], od["odom_jj"], od["Tij"], od["s_bar"], OP["sigma_odom_t"],
                    OP["sigma_odom_r"], OP["sigma_scale_prior"])
    db = sg.solve(3, 0.0)
    assert torch.equal(da, db) and torch.equal(Ta, Tb)


def test_constrain_points_to_ray_bitwise(dev):
    """geometry.py:37-42 as one kernel: bit-identical to the torch expression of the reference
    evaluated on the GPU and to numpy float32 on the CPU; the arena applies it on put_keyframe."""
    from frodo_slam_b200.graph import KeyframeGraphArena, constrain_points_to_ray
    H, W = 37, 53
    g = synth.keyframe_graph(3, H, W, [(0, 1), (1, 2)], seed=4, device=dev)
    Xs, K = g["Xs"].contiguous(), g["K"].contiguous()
    out = constrain_points_to_ray((H, W), Xs, K)
    u, v = torch.meshgrid(torch.arange(W, device=dev), torch.arange(H, device=dev), indexing="xy")
    uv = torch.stack((u, v), dim=-1).reshape(1, H * W, 2).float().repeat(3, 1, 1)
    z = Xs[..., 2:3]
    ref = torch.cat((z * ((uv[..., 0:1] - K[0, 2]) / K[0, 0]), z * ((uv[..., 1:2] - K[1, 2]) / K[1, 1]),
                     z * 1.0), dim=-1)
    assert torch.equal(out, ref)
    Kn, Xn = K.cpu().numpy(), Xs.cpu().numpy()
    un = (np.arange(H * W) % W).astype(np.float32)
    vn = (np.arange(H * W) // W).astype(np.float32)
    refn = np.stack([Xn[..., 2] * ((un - Kn[0, 2]) / Kn[0, 0]), Xn[..., 2] * ((vn - Kn[1, 2]) / Kn[1, 1]),
                     Xn[..., 2]], axis=-1)
    assert np.array_equal(out.cpu().numpy(), refn)
    A = KeyframeGraphArena(H * W, 3, 2, device=dev)
    A.put_keyframe(0, Xs[0], g["Cs"][0], g["Twc"][0], calib=((H, W), K))
    assert torch.equal(A.Xs[0], ref[0])
