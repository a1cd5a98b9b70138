"""The PyTorch-CPU formulation (oracle/torch_cpu.py, the "C1" CPU baseline of SURVEY 8d) against
the C oracle on small seeded inputs: it must be the same computation before its timing means
anything."""
import numpy as np
import torch

from helpers import GP, MP, graph_inputs, matching_inputs, oracle_gn, rel_err


def test_refine_matches_torch_equals_oracle():
    from oracle import oracle, torch_cpu
    from frodo_slam_b200 import synth
    d = synth.matching_pairs(1, 24, 32, seed=11)
    g = torch.Generator().manual_seed(11)
    p = d["p_true"].round().long() + torch.randint(-4, 5, (1, 24 * 32, 2), generator=g)
    p[:, ::37] += 30
    D11 = d["D11"].half().contiguous()
    D21 = d["D21"].half().reshape(1, 24 * 32, -1).contiguous()
    pn, sc = torch_cpu.refine_matches(D11, D21, p, MP["radius"], MP["dilation_max"])
    pr, sr = oracle.refine_matches(D11.numpy(), D21.numpy(), p.numpy(), MP["radius"],
                                   MP["dilation_max"])
    assert np.array_equal(pn.numpy(), pr)
    assert np.array_equal(sc.numpy().view(np.uint16), sr.view(np.uint16))


def test_iter_proj_torch_close_to_oracle():
    from oracle import oracle, torch_cpu
    _, rays, pts, p0 = matching_inputs(1, 40, 56, seed=5)
    p, c = torch_cpu.iter_proj(rays, pts, p0, MP["max_iter"], MP["lambda_init"],
                               MP["convergence_thresh"])
    pr, cr = oracle.iter_proj(rays.numpy(), pts.numpy(), p0.numpy(), MP["max_iter"],
                              MP["lambda_init"], MP["convergence_thresh"])
    d = np.abs(p.numpy() - pr)
    # torch evaluates the same formulas without the kernels' FMA contractions: a handful of
    # accept/reject decisions at the 1e-7 level may differ, the converged positions do not
    assert np.quantile(d, 0.99) < 1e-3 and (d > 1e-2).mean() < 0.01
    assert (c.numpy() == cr).mean() > 0.98


def test_gn_rays_torch_equals_oracle():
    from oracle import torch_cpu
    g = graph_inputs("window8", 24, 32)
    ref = oracle_gn("rays", g, 1, 1, return_system=True)
    H, gv = torch_cpu.gn_rays_system(g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                     g["valid_match"], g["Q"], GP["sigma_ray"], GP["sigma_dist"],
                                     GP["C_conf"], GP["Q_conf"], 1)
    assert rel_err(H.numpy(), ref["H"]) < 1e-4 and rel_err(gv.numpy(), ref["g"]) < 1e-4
    dx = torch_cpu.gn_solve(H, gv)
    assert rel_err(dx.numpy(), ref["dx"]) < 1e-3
