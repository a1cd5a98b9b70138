"""Device-resident keyframe / edge arena for the global optimiser.

What it replaces: ``FactorGraph`` keeps every edge as a Python list of tensors and, on EVERY
solve, stacks them (``prep_two_way_edges``, global_opt.py:196-212), pulls each keyframe's
pointmap out of CPU shared memory and stacks those too (``get_poses_points`` :214-217;
``SharedKeyframes`` frame.py:374-390) and uploads the lot -- 201 MB for 64 keyframes at
512x384, several milliseconds over PCIe, more than the whole Gauss-Newton solve takes here.

``KeyframeGraphArena`` owns fixed-capacity device arrays instead:

* keyframes: ``Twc [Kcap,8]``, ``Xs [Kcap,N,3]``, ``Cs [Kcap,N]`` and the packed copy the
  accumulate kernel streams (``frodo_gn_pack_slot``), one ring slot per keyframe id
  (``slot = id % Kcap``); a keyframe is written -- and packed -- once, when it is added or when
  the tracker updates its pointmap (tracker.py:318-320);
* edges: ``idx [2 Ecap, N]``, ``valid``, ``Q`` with both directions of an undirected edge in
  adjacent rows (the reference concatenates all forward then all backward edges,
  global_opt.py:205-210; the order only permutes the terms of the fp64 assembly sums);
* eviction (``_materialise_edges``, global_opt.py:143-176): edges touching a keyframe older
  than ``oldest_alive`` are dropped by moving the last live edge into their rows.

A solve uploads three small int32 index arrays (arena rows, system ranks, rank -> row) built
from the host-side id lists -- the reference also keeps the ids in Python lists -- and runs
``frodo_gauss_newton_arena``: no pointmap or edge tensor is touched by the host again.
No CPU fallback: tensors must live on the arena's CUDA device.
"""
import bisect
import ctypes

import torch

from . import _lib
from ._lib import OdomTerms
from .backends import make_params

__all__ = ["KeyframeGraphArena"]


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None


def constrain_points_to_ray(img_size, Xs, K, out=None):
    """mast3r_slam/geometry.py:37-42 (same name and argument order): every point keeps its depth
    and moves onto the ray of its pixel.  Xs [..., H*W, 3] f32 CUDA, K [3,3] f32 CUDA; what
    ``solve_GN_calib`` does before ``gauss_newton_calib`` (global_opt.py:296-300), one kernel."""
    h, w = int(img_size[0]), int(img_size[1])
    if not Xs.is_cuda or Xs.dtype != torch.float32 or not Xs.is_contiguous():
        raise RuntimeError("Xs must be a contiguous float32 CUDA tensor")
    if not K.is_cuda or K.dtype != torch.float32 or K.numel() != 9 or not K.is_contiguous():
        raise RuntimeError("K must be a contiguous float32 [3,3] CUDA tensor")
    if Xs.shape[-1] != 3 or Xs.numel() % (3 * h * w) != 0:
        raise RuntimeError("Xs must be [..., H*W, 3] for the given img_size")
    if out is None:
        out = torch.empty_like(Xs)
    with torch.cuda.device(Xs.device):
        _lib.check(_lib.load().frodo_constrain_points_to_ray(
            _ptr(Xs), _ptr(K), Xs.numel() // (3 * h * w), h, w, _ptr(out),
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


class KeyframeGraphArena:
    def __init__(self, num_points, max_keyframes, max_edges, device="cuda"):
        """num_points: N = H*W; max_keyframes: ring capacity (the sliding window of
        local_opt.window_size, or the whole map); max_edges: undirected edge capacity."""
        self.dev = torch.device(device)
        if self.dev.type != "cuda":
            raise RuntimeError("KeyframeGraphArena needs a CUDA device (frodo-slam_b200 has no CPU path)")
        if num_points < 1 or max_keyframes < 1 or max_edges < 1 or 2 * max_edges > 65535:
            raise ValueError("need num_points >= 1, max_keyframes >= 1, 1 <= max_edges <= 32767")
        self.N, self.Kcap, self.Ecap = int(num_points), int(max_keyframes), int(max_edges)
        self.lib = _lib.load()
        N, K, E2 = self.N, self.Kcap, 2 * self.Ecap
        with torch.cuda.device(self.dev):
            self.Twc = torch.zeros(K, 8, dtype=torch.float32, device=self.dev)
            self.Xs = torch.zeros(K, N, 3, dtype=torch.float32, device=self.dev)
            self.Cs = torch.zeros(K, N, 1, dtype=torch.float32, device=self.dev)
            self.packed = torch.zeros(self.lib.frodo_gn_packed_bytes(K, N), dtype=torch.uint8,
                                      device=self.dev)
            self.idx = torch.zeros(E2, N, dtype=torch.int64, device=self.dev)
            self.valid = torch.zeros(E2, N, 1, dtype=torch.bool, device=self.dev)
            self.Q = torch.zeros(E2, N, 1, dtype=torch.float32, device=self.dev)
        self._kf_of_slot = [None] * K   # keyframe id held by each ring slot
        self._ii, self._jj = [], []     # undirected edges (ids), row pair (2e, 2e+1) each
        self._ws = None
        self._ws_bytes = 0

    # ------------------------------------------------------------------ keyframes
    def slot_of(self, kf_id):
        s = int(kf_id) % self.Kcap
        if self._kf_of_slot[s] != int(kf_id):
            raise KeyError(f"keyframe {kf_id} is not in the arena")
        return s

    def has_keyframe(self, kf_id):
        return self._kf_of_slot[int(kf_id) % self.Kcap] == int(kf_id)

    def keyframe_ids(self):
        return sorted(k for k in self._kf_of_slot if k is not None)

    def put_keyframe(self, kf_id, X, C, T_WC=None, calib=None):
        """Add keyframe `kf_id`, or overwrite its pointmap / confidence (and pose, if given).
        X [N,3] (or [H,W,3]), C [N] (or [N,1]) f32 on the device.  The ring slot must be free
        or already hold this id: evict first (``evict_older_than``).  ``calib=(img_size, K)``:
        the pointmap is re-projected onto its pixel rays as it is stored
        (``constrain_points_to_ray``, what ``solve_GN_calib`` does per solve,
        global_opt.py:296-300), so calibrated solves run on the arena directly."""
        kf_id = int(kf_id)
        s = kf_id % self.Kcap
        if self._kf_of_slot[s] is not None and self._kf_of_slot[s] != kf_id:
            raise RuntimeError(f"ring slot {s} still holds keyframe {self._kf_of_slot[s]}; evict "
                               f"it before adding keyframe {kf_id} (capacity {self.Kcap})")
        new = self._kf_of_slot[s] is None
        if new and T_WC is None:
            raise RuntimeError("a new keyframe needs a pose")
        self._dev_f32(X, "X", self.N * 3)
        self._dev_f32(C, "C", self.N)
        with torch.cuda.device(self.dev):
            if calib is not None:
                constrain_points_to_ray(calib[0], X.reshape(self.N, 3).contiguous(), calib[1],
                                        out=self.Xs[s])
            else:
                self.Xs[s].copy_(X.reshape(self.N, 3))
            self.Cs[s].copy_(C.reshape(self.N, 1))
            if T_WC is not None:
                self.set_pose(kf_id, T_WC, _slot=s)
            _lib.check(self.lib.frodo_gn_pack_slot(
                _ptr(self.Xs[s]), _ptr(self.Cs[s]), self.Kcap, s, self.N, _ptr(self.packed),
                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        self._kf_of_slot[s] = kf_id

    def set_pose(self, kf_id, T_WC, _slot=None):
        s = self.slot_of(kf_id) if _slot is None else _slot
        d = T_WC if isinstance(T_WC, torch.Tensor) else T_WC.data
        self._dev_f32(d, "T_WC", 8)
        self.Twc[s].copy_(d.reshape(8))

    def poses(self, kf_ids):
        """[len(kf_ids), 8] poses of the given keyframes (device tensor)."""
        rows = torch.tensor([self.slot_of(k) for k in kf_ids], dtype=torch.int64, device=self.dev)
        return self.Twc.index_select(0, rows)

    # ---------------------------------------------------------------------- edges
    def num_edges(self):
        return len(self._ii)

    def edge_list(self):
        """Undirected edges (ii, jj) in arena order; directed rows are (2e: i->j data, 2e+1)."""
        return list(self._ii), list(self._jj)

    def add_edges(self, ii, jj, idx_ii2jj, idx_jj2ii, valid_match_j, valid_match_i, Q_ii2jj,
                  Q_jj2ii):
        """What ``FactorGraph.add_factors`` appends per accepted edge (global_opt.py:126-137):
        for undirected edge e = (ii[e], jj[e]) the two directed edges of
        ``prep_two_way_edges`` (:205-210): (ii -> jj: idx_ii2jj, valid_match_j, Q_ii2jj) and
        (jj -> ii: idx_jj2ii, valid_match_i, Q_jj2ii).  Tensors [e, N(,1)] on the device."""
        ii, jj = [int(x) for x in ii], [int(x) for x in jj]
        n = len(ii)
        if len(jj) != n:
            raise ValueError("ii and jj must have the same length")
        if len(self._ii) + n > self.Ecap:
            raise RuntimeError(f"edge capacity {self.Ecap} exceeded")
        for k in ii + jj:
            self.slot_of(k)
        e0 = len(self._ii)
        N = self.N
        for name, t, dt in (("idx_ii2jj", idx_ii2jj, torch.int64), ("idx_jj2ii", idx_jj2ii, torch.int64),
                            ("valid_match_j", valid_match_j, torch.bool),
                            ("valid_match_i", valid_match_i, torch.bool),
                            ("Q_ii2jj", Q_ii2jj, torch.float32), ("Q_jj2ii", Q_jj2ii, torch.float32)):
            if not t.is_cuda or t.device != self.dev or t.dtype != dt or t.numel() != n * N:
                raise RuntimeError(f"{name} must be a {dt} tensor [{n},{N}] on {self.dev}")
        rows = slice(2 * e0, 2 * (e0 + n))
        with torch.cuda.device(self.dev):
            self.idx[rows].view(n, 2, N)[:, 0].copy_(idx_ii2jj.reshape(n, N))
            self.idx[rows].view(n, 2, N)[:, 1].copy_(idx_jj2ii.reshape(n, N))
            self.valid[rows].view(n, 2, N)[:, 0].copy_(valid_match_j.reshape(n, N))
            self.valid[rows].view(n, 2, N)[:, 1].copy_(valid_match_i.reshape(n, N))
            self.Q[rows].view(n, 2, N)[:, 0].copy_(Q_ii2jj.reshape(n, N))
            self.Q[rows].view(n, 2, N)[:, 1].copy_(Q_jj2ii.reshape(n, N))
        self._ii += ii
        self._jj += jj

    def evict_older_than(self, oldest_alive):
        """``_materialise_edges`` (global_opt.py:143-176): drop every edge with an endpoint id
        < oldest_alive, then free the keyframes below it.  Returns True if edges remain."""
        oldest_alive = int(oldest_alive)
        e = 0
        with torch.cuda.device(self.dev):
            while e < len(self._ii):
                if self._ii[e] >= oldest_alive and self._jj[e] >= oldest_alive:
                    e += 1
                    continue
                last = len(self._ii) - 1
                if e != last:  # move the last live edge's two rows into the hole
                    for buf in (self.idx, self.valid, self.Q):
                        buf[2 * e:2 * e + 2].copy_(buf[2 * last:2 * last + 2])
                    self._ii[e], self._jj[e] = self._ii[last], self._jj[last]
                self._ii.pop()
                self._jj.pop()
        for s, k in enumerate(self._kf_of_slot):
            if k is not None and k < oldest_alive:
                self._kf_of_slot[s] = None
        return bool(self._ii)

    def remove_keyframe(self, kf_id):
        """Drop every edge touching `kf_id` (``FactorGraph.remove_factors_i``,
        local_mapping.py) and free its ring slot."""
        kf_id = int(kf_id)
        s = self.slot_of(kf_id)
        e = 0
        with torch.cuda.device(self.dev):
            while e < len(self._ii):
                if self._ii[e] != kf_id and self._jj[e] != kf_id:
                    e += 1
                    continue
                last = len(self._ii) - 1
                if e != last:
                    for buf in (self.idx, self.valid, self.Q):
                        buf[2 * e:2 * e + 2].copy_(buf[2 * last:2 * last + 2])
                    self._ii[e], self._jj[e] = self._ii[last], self._jj[last]
                self._ii.pop()
                self._jj.pop()
        self._kf_of_slot[s] = None

    # ---------------------------------------------------------------------- solve
    def _index_arrays(self):
        ii2 = [v for e in range(len(self._ii)) for v in (self._ii[e], self._jj[e])]
        jj2 = [v for e in range(len(self._ii)) for v in (self._jj[e], self._ii[e])]
        uniq = sorted(set(ii2))  # get_unique_kf_idx, global_opt.py:180-184
        rank = {k: r for r, k in enumerate(uniq)}
        flat = ([self.slot_of(k) for k in ii2] + [self.slot_of(k) for k in jj2] +
                [rank[k] for k in ii2] + [rank[k] for k in jj2] + [self.slot_of(k) for k in uniq])
        t = torch.tensor(flat, dtype=torch.int32).to(self.dev, non_blocking=True)
        E = len(ii2)
        return uniq, t[0:E], t[E:2 * E], t[2 * E:3 * E], t[3 * E:4 * E], t[4 * E:], E

    def _odom_terms(self, uniq, odom):
        """frodo_gn_odom_terms for `odom` = dict(odom_ii, odom_jj: keyframe ids (lists or
        tensors), Tij [Eo,7|8] f32 device, s_bar [Ku] by rank of the unique keyframes or None,
        sigma_odom_t, sigma_odom_r, sigma_scale_prior).  Rows of Twc = arena slots; system
        ranks = searchsorted in the visual keyframe list (gn_kernels.cu:1829)."""
        oi = [int(x) for x in (odom["odom_ii"].tolist() if torch.is_tensor(odom["odom_ii"])
                               else odom["odom_ii"])]
        oj = [int(x) for x in (odom["odom_jj"].tolist() if torch.is_tensor(odom["odom_jj"])
                               else odom["odom_jj"])]
        Eo = len(oi)
        if len(oj) != Eo:
            raise ValueError("odom_ii and odom_jj must have the same length")
        rows = torch.tensor([self.slot_of(k) for k in oi] + [self.slot_of(k) for k in oj],
                            dtype=torch.int64).to(self.dev, non_blocking=True)
        rank = torch.tensor([bisect.bisect_left(uniq, k) for k in oi] +
                            [bisect.bisect_left(uniq, k) for k in oj],
                            dtype=torch.int32).to(self.dev, non_blocking=True)
        Tij = odom["Tij"]
        if Eo > 0 and (not Tij.is_cuda or Tij.dtype != torch.float32 or Tij.dim() != 2 or
                       Tij.shape[0] != Eo or Tij.shape[1] < 7 or not Tij.is_contiguous()):
            raise RuntimeError("Tij must be a contiguous float32 [num_odom_edges, 7 or 8] device tensor")
        sb = odom.get("s_bar")
        if sb is not None:
            sb = torch.as_tensor(sb).to(device=self.dev, dtype=torch.float64).contiguous()
            if sb.numel() != len(uniq):
                raise RuntimeError("s_bar must hold one scale target per participating keyframe")
        nb = self.lib.frodo_gn_odom_blocks_bytes(Eo)
        blocks = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        t = OdomTerms()
        t.odom_ii_row = rows[:Eo].data_ptr() if Eo else None
        t.odom_jj_row = rows[Eo:].data_ptr() if Eo else None
        t.odom_rank = rank.data_ptr() if Eo else None
        t.Tij = Tij.data_ptr() if Eo else None
        t.tij_stride = int(Tij.shape[1]) if Eo else 0
        t.E_odom = Eo
        t.s_bar = sb.data_ptr() if sb is not None else None
        t.sigma_odom_t = float(odom["sigma_odom_t"])
        t.sigma_odom_r = float(odom["sigma_odom_r"])
        t.sigma_scale_prior = float(odom["sigma_scale_prior"])
        t.blocks = blocks.data_ptr()
        t.blocks_bytes = nb
        return t, (rows, rank, sb, blocks, Tij)  # keep the tensors alive through the launch

    def solve(self, mode, num_fix, max_iter, delta_thresh, K=None, odom=None, **params):
        """One global optimisation over the edges currently in the arena, poses updated in
        place.  mode / params as ``backends.make_params`` ("rays": sigma_ray, sigma_dist,
        C_thresh, Q_thresh; "points": sigma_point, ...; "calib": + height, width,
        pixel_border, z_eps and K [3,3]).  Returns (dx [Ku - num_fix, 7], unique keyframe ids):
        the same dx the drop-in operator returns for the stacked inputs.  With odometry terms
        (``odom``) that equality needs keyframe id == row of the stacked ``Twc`` (ids 0..K-1 in
        order): the drop-in op and the reference index ``Twc`` by the RAW odometry ids
        (gn_kernels.cu:1121-1122) while the arena looks the ids up in its slots."""
        if not self._ii:
            raise RuntimeError("no edges in the arena")
        prm = make_params(mode, **params)
        uniq, ii_slot, jj_slot, ii_rank, jj_rank, pose_slot, E = self._index_arrays()
        Ku = len(uniq)
        if mode == "calib":
            if K is None or not K.is_cuda or K.dtype != torch.float32 or K.numel() != 9:
                raise RuntimeError("calib needs K [3,3] float32 on the device")
            K = K.contiguous()
        if odom is not None and mode != "rays":
            raise ValueError("odometry terms exist for the rays residual only")
        with torch.cuda.device(self.dev):
            dx = torch.zeros((max(Ku - num_fix, 0), 7), dtype=torch.float32, device=self.dev)
            if max_iter <= 0 or Ku - num_fix <= 0:
                return dx, uniq
            terms, keep = self._odom_terms(uniq, odom) if odom is not None else (None, None)
            nbytes = self.lib.frodo_gn_arena_workspace_bytes(self.Kcap, E, self.N)
            if self._ws is None or self._ws_bytes < nbytes:
                self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
                self._ws_bytes = nbytes
            _lib.check(self.lib.frodo_gauss_newton_arena(
                ctypes.byref(prm), _ptr(self.Twc), _ptr(self.Xs), _ptr(self.Cs), _ptr(K),
                _ptr(self.packed), _ptr(ii_slot), _ptr(jj_slot), _ptr(ii_rank), _ptr(jj_rank),
                _ptr(pose_slot), _ptr(self.idx), _ptr(self.valid), _ptr(self.Q), self.Kcap, Ku,
                self.N, E, int(num_fix), int(max_iter), float(delta_thresh),
                ctypes.byref(terms) if terms is not None else None, _ptr(dx), None,
                _ptr(self._ws), self._ws_bytes,
                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
            del keep
        return dx, uniq

    def solve_rays_odom(self, odom_ii, odom_jj, Tij, s_bar, sigma_odom_t, sigma_odom_r, sigma_ray,
                        sigma_dist, sigma_scale_prior, C_thresh, Q_thresh, num_fix, max_iter,
                        delta_thresh):
        """The arena form of ``gauss_newton_rays_odom`` (gn.cpp:29-45; the op the fork's
        ``solve_GN_rays`` calls, global_opt.py:260): same scalar arguments in the same order;
        odom_ii / odom_jj are keyframe ids, s_bar [number of participating keyframes]."""
        odom = dict(odom_ii=odom_ii, odom_jj=odom_jj, Tij=Tij, s_bar=s_bar,
                    sigma_odom_t=sigma_odom_t, sigma_odom_r=sigma_odom_r,
                    sigma_scale_prior=sigma_scale_prior)
        return self.solve("rays", num_fix, max_iter, delta_thresh, odom=odom, sigma_ray=sigma_ray,
                          sigma_dist=sigma_dist, C_thresh=C_thresh, Q_thresh=Q_thresh)

    def solve_rays(self, sigma_ray, sigma_dist, C_thresh, Q_thresh, num_fix, max_iter,
                   delta_thresh):
        """The arena form of ``gauss_newton_rays`` (same scalar arguments, gn.cpp:63-74)."""
        return self.solve("rays", num_fix, max_iter, delta_thresh, sigma_ray=sigma_ray,
                          sigma_dist=sigma_dist, C_thresh=C_thresh, Q_thresh=Q_thresh)

    def stacked_inputs(self):
        """The tensors the reference would build for the same graph (``get_poses_points`` +
        ``prep_two_way_edges`` in arena edge order): for tests and for callers that still want
        the drop-in operator.  Returns (uniq, Twc, Xs, Cs, ii, jj, idx, valid, Q)."""
        uniq, ii_slot, jj_slot, ii_rank, jj_rank, pose_slot, E = self._index_arrays()
        rows = pose_slot.long()
        ids = torch.tensor(uniq, dtype=torch.int64, device=self.dev)
        return (uniq, self.Twc[rows].contiguous(), self.Xs[rows].contiguous(),
                self.Cs[rows].contiguous(), ids[ii_rank.long()], ids[jj_rank.long()],
                self.idx[:E].contiguous(), self.valid[:E].contiguous(), self.Q[:E].contiguous())

    def _dev_f32(self, t, name, numel):
        if not isinstance(t, torch.Tensor) or not t.is_cuda or t.device != self.dev:
            raise RuntimeError(f"{name} must be a CUDA tensor on {self.dev} (no CPU path)")
        if t.dtype != torch.float32 or t.numel() != numel:
            raise RuntimeError(f"{name} must be float32 with {numel} values")
