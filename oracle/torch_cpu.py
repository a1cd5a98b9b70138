"""PyTorch-CPU formulation of the hot path (SURVEY 8(d) "C1", BASELINE.json configs[0]).

TEST / BENCH INFRASTRUCTURE, not product code: a vectorised torch restatement of iter_proj,
refine_matches and one gauss_newton_rays iteration, run on the host cores with
``torch.set_num_threads(os.cpu_count())``.  It is what a PyTorch user would write without the
CUDA extension -- the reference itself has no PyTorch version of iter_proj / refine_matches; its
only PyTorch Gauss-Newton is the tracker's single-pair solve (tracker.py:397-465,
geometry.py:17-104, nonlinear_optimizer.py:28-33), whose whitening / Huber / H = A^T A
structure ``gn_rays_iteration`` follows.  Formula sources: matching_kernels.cu:45-84 (refine),
:148-282 (iter_proj), gn_kernels.cu:673-861 (ray residual + Jacobian), :214-339 (Sim3 helpers),
:72-122, :163-199 (assembly, solve).  Only tests/ and bench.py's cpu_baseline legs import it.

Checked against the C oracle in tests/test_torch_cpu.py: refine_matches bit-exact, iter_proj to
1e-3 px (torch does not contract to FMA the way the kernels do), H, g and dx to 1e-4."""
import os

import torch


def use_all_cores():
    n = os.cpu_count() or 1
    torch.set_num_threads(n)
    return n


# --------------------------------------------------------------------------- iter_proj

def _bilinear(img, u, v, W, nch):
    """img [H*W, 9]; returns the first `nch` channels blended at (u, v) with the reference's
    (crossed) weight naming, matching_kernels.cu:164-192."""
    u11 = torch.floor(u)
    v11 = torch.floor(v)
    du, dv = u - u11, v - v11
    iu, iv = u11.long(), v11.long()
    w11 = du * dv
    w12 = (1.0 - du) * dv
    w21 = du * (1.0 - dv)
    w22 = (1.0 - du) * (1.0 - dv)
    base = iv * W + iu
    r11 = img[base + W + 1, :nch]
    r12 = img[base + W, :nch]
    r21 = img[base + 1, :nch]
    r22 = img[base, :nch]
    return w11[:, None] * r11 + w12[:, None] * r12 + w21[:, None] * r21 + w22[:, None] * r22


def iter_proj(rays_img, pts, p_init, max_iter, lambda_init, cost_thresh):
    """rays_img [B,H,W,9], pts [B,N,3], p_init [B,N,2] (float32 CPU tensors) ->
    p_new [B,N,2], converged [B,N] bool."""
    B, H, W, _ = rays_img.shape
    N = pts.shape[1]
    p_new = torch.empty_like(p_init)
    conv_out = torch.zeros(B, N, dtype=torch.bool)
    for b in range(B):
        img = rays_img[b].reshape(H * W, 9)
        t = pts[b]
        u = p_init[b, :, 0].clamp(1.0, float(W - 2))
        v = p_init[b, :, 1].clamp(1.0, float(H - 2))
        lam = torch.full((N,), float(lambda_init), dtype=torch.float64)
        conv = torch.zeros(N, dtype=torch.bool)
        for _ in range(max_iter):
            f = _bilinear(img, u, v, W, 9)
            r, gx, gy = f[:, 0:3], f[:, 3:6], f[:, 6:9]
            r = r / r.norm(dim=1, keepdim=True)
            err = r - t
            cost = (err * err).sum(1)
            lamf = lam.float()
            A00 = (gx * gx).sum(1) + lamf
            A01 = (gx * gy).sum(1)
            A11 = (gy * gy).sum(1) + lamf
            b0 = -(err * gx).sum(1)
            b1 = -(err * gy).sum(1)
            det_inv = 1.0 / (A00 * A11 - A01 * A01)
            u_new = (u + det_inv * (A11 * b0 - A01 * b1)).clamp(1.0, float(W - 2))
            v_new = (v + det_inv * (-A01 * b0 + A00 * b1)).clamp(1.0, float(H - 2))
            r2 = _bilinear(img, u_new, v_new, W, 3)
            r2 = r2 / r2.norm(dim=1, keepdim=True)
            e2 = r2 - t
            new_cost = (e2 * e2).sum(1)
            acc = new_cost < cost
            u = torch.where(acc, u_new, u)
            v = torch.where(acc, v_new, v)
            lam = torch.where(acc, lam * 0.1, lam * 10.0).float().double()
            conv = torch.where(acc, new_cost < cost_thresh, cost < cost_thresh)
        p_new[b, :, 0] = u
        p_new[b, :, 1] = v
        conv_out[b] = conv
    return p_new, conv_out


# ----------------------------------------------------------------------- refine_matches

def refine_matches(D11, D21, p1, radius, dilation_max):
    """D11 [B,H,W,F], D21 [B,N,F] float16, p1 [B,N,2] int64 (u, v) -> p1_new, max_scores (fp16).
    Every multiply and every add rounds to fp16 (c10::Half), k ascending, candidates in the
    reference's order (d descending, u outer, v inner), strict '>' from a start of 0."""
    B, H, W, F = D11.shape
    N = D21.shape[1]
    p_out = torch.empty_like(p1)
    s_out = torch.empty(B, N, dtype=torch.float16)
    for b in range(B):
        img = D11[b].reshape(H * W, F)
        q = D21[b]
        u0, v0 = p1[b, :, 0].clone(), p1[b, :, 1].clone()
        best = torch.zeros(N, dtype=torch.float16)
        bu, bv = u0.clone(), v0.clone()
        for d in range(dilation_max, 0, -1):
            rd = radius * d
            for i in range(0, 2 * rd + 1, d):
                for j in range(0, 2 * rd + 1, d):
                    u = u0 - rd + i
                    v = v0 - rd + j
                    inside = (v >= 0) & (v < H) & (u >= 0) & (u < W)
                    c = img[(v.clamp(0, H - 1) * W + u.clamp(0, W - 1))]
                    score = torch.zeros(N, dtype=torch.float16)
                    for k in range(F):
                        score = score + q[:, k] * c[:, k]  # two fp16 roundings
                    better = inside & (score.float() > best.float())
                    best = torch.where(better, score, best)
                    bu = torch.where(better, u, bu)
                    bv = torch.where(better, v, bv)
            u0, v0 = bu.clone(), bv.clone()
        p_out[b, :, 0] = bu
        p_out[b, :, 1] = bv
        s_out[b] = best
    return p_out, s_out


# ------------------------------------------------------------------- gauss_newton_rays

def _act_so3(q, X):
    """q [4] xyzw, X [...,3]."""
    qv, w = q[:3], q[3]
    uv = 2.0 * torch.cross(qv.expand_as(X), X, dim=-1)
    return X + w * uv + torch.cross(qv.expand_as(X), uv, dim=-1)


def _quat_mul(a, b):
    return torch.stack([a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1],
                        a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0],
                        a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3],
                        a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]])


def _rel_sim3(Ti, Tj):
    si_inv = 1.0 / Ti[7]
    qi_inv = Ti[3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0])
    return _act_so3(qi_inv, Tj[0:3] - Ti[0:3]) * si_inv, _quat_mul(qi_inv, Tj[3:7]), si_inv * Tj[7]


def _adj_inv_matrix(T):
    """7x7 M with apply_Sim3_adj_inv(T, X) = M X (gn_kernels.cu:319-339): the map from the local
    tangent [I, -[P]x, P] to the pose-j tangent, constant over the points of an edge."""
    return torch.stack([_apply_adj_inv(T, e) for e in torch.eye(7)], dim=1)


def _apply_adj_inv(T, X):
    t, q, s_inv = T[0:3], T[3:7], 1.0 / T[7]
    Ra = _act_so3(q, X[0:3])
    out = torch.empty(7)
    out[0:3] = s_inv * Ra
    out[3:6] = _act_so3(q, X[3:6]) + s_inv * torch.cross(t, Ra, dim=0)
    out[6] = X[6] + s_inv * (t * Ra).sum()
    return out


def _huber(r, k=1.345):
    a = r.abs()
    return torch.where(a < k, torch.ones_like(a), k / a)


def edge_system_rays(Ti, Tj, Xi_all, Xj, Ci_all, Cj, idx, valid_match, Q, sigma_ray, sigma_dist,
                     C_thresh, Q_thresh):
    """One directed edge: whitened 4-row residual b [N,4] and Jacobian wrt pose j, A [N,4,7]
    (J_i = -J_j), then H_jj = A^T W A, g_j = A^T W b as tracker.py:397-412 forms them."""
    tij, qij, sij = _rel_sim3(Ti, Tj)
    ind = torch.where(valid_match, idx, torch.zeros_like(idx))
    Xi = Xi_all[ind]
    P = sij * _act_so3(qij, Xj) + tij
    ni = Xi.norm(dim=1, keepdim=True)
    nj = P.norm(dim=1, keepdim=True)
    ri, rj = Xi / ni, P / nj
    err = torch.cat([rj - ri, nj - ni], dim=1)                      # [N,4]
    valid = valid_match & (Q > Q_thresh) & (Ci_all[ind] > C_thresh) & (Cj > C_thresh)
    sq = torch.sqrt(Q) * valid
    sw = torch.stack([sq / sigma_ray] * 3 + [sq / sigma_dist], dim=1)  # [N,4] whitening
    w = _huber(sw * err) * sw * sw
    # d(ray)/dP = (I - r r^T)/|P|, local tangent [I, -[P]x, P]  (geometry.py:17-52)
    N = P.shape[0]
    eye = torch.eye(3).expand(N, 3, 3)
    drdP = (eye - rj[:, :, None] * rj[:, None, :]) / nj[:, :, None]
    Px = torch.zeros(N, 3, 3)
    Px[:, 0, 1], Px[:, 0, 2] = -P[:, 2], P[:, 1]
    Px[:, 1, 0], Px[:, 1, 2] = P[:, 2], -P[:, 0]
    Px[:, 2, 0], Px[:, 2, 1] = -P[:, 1], P[:, 0]
    dPdx = torch.cat([eye, -Px, P[:, :, None]], dim=2)               # [N,3,7]
    Jx = torch.cat([drdP @ dPdx, (rj[:, None, :] @ dPdx)], dim=1)     # [N,4,7]
    M = _adj_inv_matrix(Ti)
    A = Jx @ M.T                                                      # pose-j tangent
    WA = w[:, :, None] * A
    H = torch.einsum("nri,nrj->ij", WA.double(), A.double())
    g = torch.einsum("nri,nr->i", WA.double(), err.double())
    return H, g


def gn_rays_system(Twc, Xs, Cs, ii, jj, idx, valid_match, Q, sigma_ray, sigma_dist, C_thresh,
                   Q_thresh, num_fix, edges=None):
    """Dense normal equations (H [n,n], g [n], fp64) of one gauss_newton_rays iteration
    (gn_kernels.cu:72-122 semantics: blocks touching a fixed keyframe are skipped).
    `edges`: subset of edge indices to accumulate (bounded CPU samples)."""
    ids = torch.unique(torch.cat([ii, jj]), sorted=True)
    ie = torch.searchsorted(ids, ii)
    je = torch.searchsorted(ids, jj)
    K, N = Xs.shape[0], Xs.shape[1]
    n = 7 * (K - num_fix)
    H = torch.zeros(n, n, dtype=torch.float64)
    g = torch.zeros(n, dtype=torch.float64)
    Cs2 = Cs.reshape(K, N)
    for e in (range(ii.shape[0]) if edges is None else edges):
        i, j = int(ie[e]), int(je[e])
        A, gj = edge_system_rays(Twc[i], Twc[j], Xs[i], Xs[j], Cs2[i], Cs2[j], idx[e].reshape(N),
                                 valid_match[e].reshape(N), Q[e].reshape(N), sigma_ray, sigma_dist,
                                 C_thresh, Q_thresh)
        io, jo = i - num_fix, j - num_fix
        si, sj = slice(7 * io, 7 * io + 7), slice(7 * jo, 7 * jo + 7)
        if io >= 0:
            H[si, si] += A
            g[si] -= gj
        if jo >= 0:
            H[sj, sj] += A
            g[sj] += gj
        if io >= 0 and jo >= 0:
            H[si, sj] -= A
            H[sj, si] -= A
    return H, g


def gn_solve(H, g):
    """dx = -H^-1 g by Cholesky (zeros on failure, gn_kernels.cu:163-199), float32 [K-fix,7]."""
    L, info = torch.linalg.cholesky_ex(H)
    if int(info) != 0:
        return torch.zeros(H.shape[0] // 7, 7)
    return (-torch.cholesky_solve(g[:, None], L)[:, 0]).float().reshape(-1, 7)
