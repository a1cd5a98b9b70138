"""Edge-sharded multi-GPU Gauss-Newton: one process per GPU, keyframe
poses/pointmaps replicated, the directed edges split into contiguous per-rank slices.

Fast path (NVLink peer memory available): the whole solve is ONE C-ABI call,
``frodo_gauss_newton_sharded``.  The exchange step is fused into the accumulate kernel: the last
CTA of every edge stores the edge's 35 fp64 values into every rank's copy of the global per-edge
block array and bumps that rank's arrival counter; the Cholesky's tile build waits for the
counter and assembles in fixed edge order, so all ranks hold bit-identical systems, run the
identical solve and keep bit-identical poses with no broadcast of dx.

Fallback (no peer mapping, ``FRODO_PEER_REDUCE=0``, gloo on CPU-only test runs of the host
logic): every rank assembles its edges into a dense fp64 [H | g], one all-reduce (NCCL / gloo)
sums them, every rank solves.  The reference has no multi-GPU path."""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib
from ._lib import GnPeer, OdomTerms
from .backends import _gn_validate, _ptr, _stream


# Peer-memory exchange state, shared by every ShardedGaussNewton of a process (one symmetric
# allocation per (group, size); the epoch counter must advance identically on all ranks, which it
# does because all ranks execute the same sequence of solves).
_PEER_CACHE = {}


def _peer_state(numel, dev, group):
    """Symmetric (peer-mapped) double buffer for `numel` doubles + signal words, or None when
    torch's symmetric memory cannot serve this group (then NCCL's all-reduce is used).
    FRODO_PEER_REDUCE=0 forces NCCL."""
    if os.environ.get("FRODO_PEER_REDUCE", "1") == "0" or dev.type != "cuda":
        return None
    key = (_group_key(group), dev.index, numel)
    if key in _PEER_CACHE:
        return _PEER_CACHE[key]
    st = None
    try:
        import torch.distributed._symmetric_memory as symm
        pg = group if group is not None else dist.group.WORLD
        world = dist.get_world_size(pg)
        if world <= 16:
            padded = (numel + 1) // 2 * 2
            buf = symm.empty(2 * padded, dtype=torch.float64, device=dev)
            buf.zero_()
            hdl = symm.rendezvous(buf, pg)
            words = int(hdl.signal_pad_size) // 4
            st = dict(buf=buf, hdl=hdl, bufs_dev=int(hdl.buffer_ptrs_dev),
                      pads_dev=int(hdl.signal_pad_ptrs_dev), stride=padded * 8,
                      slot0=words - 64 if words >= 128 else 0, epoch=0,
                      rank=dist.get_rank(pg), world=world)
            torch.cuda.synchronize(dev)
            dist.barrier(pg)
    except Exception as exc:  # no peer mapping on this system / build: NCCL serves the exchange
        st = None
        if os.environ.get("FRODO_PEER_REDUCE_DEBUG"):
            print(f"[frodo] peer all-reduce unavailable, using NCCL: {exc!r}", flush=True)
    _PEER_CACHE[key] = st
    return st


_BLOCK_CACHE = {}


def _group_key(group):
    pg = group if group is not None else dist.group.WORLD
    try:
        return (pg.group_name, tuple(dist.get_process_group_ranks(pg)))
    except Exception:
        return (id(pg),)


def _block_exchange_state(max_edges, dev, group):
    """Peer-mapped per-edge block arrays + arrival counters for ``frodo_gauss_newton_sharded``
    (layout: include/frodo_b200.h, struct frodo_gn_peer), or None when symmetric memory cannot
    serve this group.  Whether the fast path is used is agreed by ALL ranks (a MIN all-reduce of
    the per-rank outcome), so one rank can never spin on peers that fell back to NCCL."""
    if dev.type != "cuda":
        return None
    key = (_group_key(group), dev.index, int(max_edges))
    if key in _BLOCK_CACHE:
        return _BLOCK_CACHE[key]
    pg = group if group is not None else dist.group.WORLD
    world = dist.get_world_size(pg)
    st = None
    want = os.environ.get("FRODO_PEER_REDUCE", "1") != "0" and world <= 16
    if want:
        try:
            import torch.distributed._symmetric_memory as symm
            nbytes = _lib.load().frodo_gn_peer_buffer_bytes(int(max_edges))
            buf = symm.empty(nbytes // 8, dtype=torch.float64, device=dev)
            buf.zero_()
            hdl = symm.rendezvous(buf, pg)
            st = dict(buf=buf, hdl=hdl, bufs_dev=int(hdl.buffer_ptrs_dev), epoch=0,
                      arrivals=[0, 0], rank=dist.get_rank(pg), world=world,
                      max_edges=int(max_edges))
            torch.cuda.synchronize(dev)
        except Exception as exc:  # no peer mapping on this system / build
            st = None
            if os.environ.get("FRODO_PEER_REDUCE_DEBUG"):
                print(f"[frodo] block exchange unavailable, using all-reduce: {exc!r}", flush=True)
    ok = torch.tensor([1 if st is not None else 0], device=dev, dtype=torch.int32)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=pg)  # also the barrier after zeroing
    if int(ok.item()) == 0:
        st = None
    _BLOCK_CACHE[key] = st
    return st


def edge_slice(num_edges, rank, world):
    """Contiguous balanced slice [lo, hi) of the edge list owned by `rank`."""
    base, rem = divmod(num_edges, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class ShardedGaussNewton:
    """Holds the per-rank buffers of one keyframe graph so repeated solves (and CUDA-event
    timing) do not reallocate.  `params` is a backends.GnParams; tensors are this rank's
    device copies: Twc/Xs/Cs(/K) replicated, ii/jj FULL edge lists (the id ranking must see
    every keyframe), idx/valid/Q only the local slice [lo, hi)."""

    def __init__(self, params, Twc, Xs, Cs, K, ii, jj, idx_local, valid_local, Q_local, lo, hi,
                 num_fix, group=None):
        self.params, self.group = params, group
        self.Twc, self.Xs, self.Cs, self.K = Twc, Xs, Cs, K
        self.lo, self.hi, self.num_fix = lo, hi, num_fix
        self.idx, self.valid, self.Q = idx_local, valid_local, Q_local
        dev = Twc.device
        self.num_kf, self.N = Xs.shape[0], Xs.shape[1]
        E = ii.shape[0]
        self.E_local = hi - lo
        lib = self.lib = _lib.load()
        n = self.n = 7 * (self.num_kf - num_fix)
        with torch.cuda.device(dev):
            self.nbytes = lib.frodo_gn_workspace_bytes(self.num_kf, max(self.E_local, 1), self.N)
            self.ws = torch.empty(self.nbytes, dtype=torch.uint8, device=dev)
            self.Hg = torch.empty((n + 1) * n, dtype=torch.float64, device=dev)
            self.dx = torch.zeros((self.num_kf - num_fix, 7), dtype=torch.float32, device=dev)
            self.flags = torch.zeros(4, dtype=torch.int32, device=dev)  # [done, iters, fails, -]
            ii_row = torch.empty(E, dtype=torch.int32, device=dev)
            jj_row = torch.empty(E, dtype=torch.int32, device=dev)
            _lib.check(lib.frodo_gn_edge_index(_ptr(ii), _ptr(jj), E, _ptr(ii_row), _ptr(jj_row),
                                               None, _stream()))
            self._ids = torch.unique(torch.cat([ii, jj]), sorted=True)  # for set_odometry
            # the exchange step: one kernel over NVLink peer memory (csrc/peer_reduce.cu) when
            # torch can map the ranks' buffers into each other, NCCL's all-reduce otherwise
            # exchange step: fused into the accumulate kernel over NVLink peer memory
            # (frodo_gauss_newton_sharded) when every rank can map its peers' buffers; otherwise
            # a dense all-reduce (peer kernel of csrc/peer_reduce.cu or NCCL / gloo)
            self.peer = None
            self.blocks = None
            self.E = E
            self.ii_row_all, self.jj_row_all = ii_row, jj_row
            if dist.is_initialized() and dist.get_world_size(group) > 1 and n > 0:
                if os.environ.get("FRODO_BLOCK_EXCHANGE", "1") != "0":
                    self.blocks = _block_exchange_state(E, dev, group)
                if self.blocks is None:
                    self.peer = _peer_state((n + 1) * n, dev, group)
            # workspace of the one-call driver: sized for all E edges
            self.nbytes_all = lib.frodo_gn_workspace_bytes(self.num_kf, max(E, 1), self.N)
            self.ws_all = None
            self.odom = None
            self.ii_row = ii_row[lo:hi].contiguous()
            self.jj_row = jj_row[lo:hi].contiguous()
            # packed fast path: pointmaps repacked inside the workspace, once per solve
            self.xc = None
            if self.N % 16 == 0 and self.E_local > 0:
                off = lib.frodo_gn_workspace_packed_offset(self.num_kf, max(self.E_local, 1), self.N)
                self.xc = ctypes.c_void_p(self.ws.data_ptr() + off)

    def set_odometry(self, odom_ii, odom_jj, Tij, s_bar, sigma_odom_t, sigma_odom_r,
                     sigma_scale_prior):
        """The extra terms of gauss_newton_rays_odom (gn_kernels.cu:1792-1934): every rank adds
        them to the summed system after the all-reduce (replicated: O(K) work, identical bits on
        every rank).  Same conventions as the single-GPU op: ids index Twc rows directly and
        are ranked against the visual keyframe list (:1121-1122, :1829)."""
        dev = self.Twc.device
        Eo = int(odom_ii.shape[0])
        rank = torch.cat([torch.searchsorted(self._ids, odom_ii.to(dev)),
                          torch.searchsorted(self._ids, odom_jj.to(dev))]).to(torch.int32)
        sb = None
        if s_bar is not None:
            sb = torch.as_tensor(s_bar).to(device=dev, dtype=torch.float64).contiguous()
        nb = self.lib.frodo_gn_odom_blocks_bytes(Eo)
        blocks = torch.empty(nb, dtype=torch.uint8, device=dev)
        oi, oj = odom_ii.to(dev).contiguous(), odom_jj.to(dev).contiguous()
        Tij = Tij.to(dev).contiguous()
        t = OdomTerms()
        t.odom_ii_row = oi.data_ptr() if Eo else None
        t.odom_jj_row = oj.data_ptr() if Eo else None
        t.odom_rank = rank.data_ptr() if Eo else None
        t.Tij = Tij.data_ptr() if Eo else None
        t.tij_stride = int(Tij.shape[1]) if Eo else 0
        t.E_odom = Eo
        t.s_bar = sb.data_ptr() if sb is not None else None
        t.sigma_odom_t, t.sigma_odom_r = float(sigma_odom_t), float(sigma_odom_r)
        t.sigma_scale_prior = float(sigma_scale_prior)
        t.blocks, t.blocks_bytes = blocks.data_ptr(), nb
        self.odom = (t, (oi, oj, rank, Tij, sb, blocks))

    def _solve_one_call(self, max_iter, delta_thresh):
        """frodo_gauss_newton_sharded: the whole loop in one C-ABI call (world 1: no exchange)."""
        lib, dev = self.lib, self.Twc.device
        with torch.cuda.device(dev):
            if self.ws_all is None:
                self.ws_all = torch.empty(self.nbytes_all, dtype=torch.uint8, device=dev)
            self.flags.zero_()
            peer = None
            b = self.blocks
            if b is not None:
                peer = GnPeer()
                peer.peer_buffers = b["bufs_dev"]
                peer.local_buffer = b["buf"].data_ptr()
                peer.world, peer.rank, peer.max_edges = b["world"], b["rank"], b["max_edges"]
                peer.epoch0 = b["epoch"] & 0xFFFFFFFF
                peer.arrivals0[0] = b["arrivals"][0] & 0xFFFFFFFF
                peer.arrivals0[1] = b["arrivals"][1] & 0xFFFFFFFF
            _lib.check(lib.frodo_gauss_newton_sharded(
                ctypes.byref(self.params), _ptr(self.Twc), _ptr(self.Xs), _ptr(self.Cs),
                _ptr(self.K), _ptr(self.ii_row_all), _ptr(self.jj_row_all), self.E, self.lo,
                self.hi, _ptr(self.idx), _ptr(self.valid), _ptr(self.Q), self.num_kf, self.N,
                self.num_fix, int(max_iter), float(delta_thresh),
                ctypes.byref(peer) if peer is not None else None,
                ctypes.byref(self.odom[0]) if self.odom is not None else None, _ptr(self.dx),
                ctypes.c_void_p(self.flags.data_ptr()), ctypes.c_void_p(self.flags.data_ptr() + 4),
                _ptr(self.ws_all), self.nbytes_all, _stream()))
            if b is not None:  # advance the exchange state exactly as the C side did
                for it in range(int(max_iter)):
                    b["epoch"] += 1
                    b["arrivals"][b["epoch"] & 1] += self.E
        return self.dx

    def solve(self, max_iter, delta_thresh):
        multi = dist.is_initialized() and dist.get_world_size(self.group) > 1
        if self.blocks is not None or (not multi and os.environ.get("FRODO_ONE_CALL", "1") != "0"):
            return self._solve_one_call(max_iter, delta_thresh)
        lib, dev = self.lib, self.Twc.device
        with torch.cuda.device(dev):
            st = _stream()
            self.flags.zero_()
            _lib.check(lib.frodo_gn_workspace_init(_ptr(self.ws), self.nbytes, self.num_kf,
                                                   max(self.E_local, 1), self.N, st))
            done = ctypes.c_void_p(self.flags.data_ptr())
            info = ctypes.c_void_p(self.flags.data_ptr() + 4)
            if self.xc is not None and max_iter > 0:
                _lib.check(lib.frodo_gn_pack_points(_ptr(self.Xs), _ptr(self.Cs), self.num_kf,
                                                    self.N, self.xc, st))
            pr = self.peer
            multi = dist.is_initialized() and dist.get_world_size(self.group) > 1
            for _ in range(max_iter):
                hg_local = _ptr(self.Hg)
                if pr is not None:  # this rank's partial goes to the symmetric buffer of this epoch
                    pr["epoch"] += 1
                    off = (pr["epoch"] & 1) * pr["stride"]
                    hg_local = ctypes.c_void_p(pr["buf"].data_ptr() + off)
                _lib.check(lib.frodo_gn_accumulate(
                    ctypes.byref(self.params), _ptr(self.Twc), _ptr(self.Xs), _ptr(self.Cs),
                    _ptr(self.K), self.xc, _ptr(self.ii_row), _ptr(self.jj_row), _ptr(self.idx),
                    _ptr(self.valid), _ptr(self.Q), self.num_kf, self.N, self.E_local,
                    self.num_fix, hg_local, None, done, _ptr(self.ws), self.nbytes, st))
                if pr is not None:
                    # never skipped (done = NULL): the epochs stay consecutive on every rank,
                    # which the double-buffering argument of peer_reduce.cu relies on
                    _lib.check(lib.frodo_gn_peer_allreduce(
                        ctypes.c_void_p(pr["bufs_dev"]), ctypes.c_void_p(pr["pads_dev"]), off,
                        pr["slot0"], pr["rank"], pr["world"], pr["epoch"] & 0xFFFFFFFF or 1,
                        _ptr(self.Hg), (self.n + 1) * self.n, None, st))
                elif multi:
                    dist.all_reduce(self.Hg, op=dist.ReduceOp.SUM, group=self.group)
                if self.odom is not None:
                    _lib.check(lib.frodo_gn_add_odom(ctypes.byref(self.odom[0]), _ptr(self.Twc),
                                                     None, self.num_kf, self.num_fix,
                                                     _ptr(self.Hg), done, st))
                _lib.check(lib.frodo_gn_solve_retract(
                    _ptr(self.Hg), _ptr(self.Twc), _ptr(self.dx), self.num_kf, self.num_fix,
                    float(delta_thresh), done, info, _ptr(self.ws), self.nbytes, st))
        return self.dx


def gauss_newton_sharded(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, num_fix,
                         max_iter, delta_thresh, group=None):
    """Multi-GPU counterpart of gauss_newton_{rays,points,calib}: every rank passes the FULL
    edge tensors (as the single-GPU op takes them) and works on its own slice.  Updates Twc in
    place on every rank, returns [dx]."""
    _gn_validate(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, K)
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    E, N = ii.shape[0], Xs.shape[1]
    lo, hi = edge_slice(E, rank, world)
    s = ShardedGaussNewton(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj.view(E, N)[lo:hi],
                           valid_match.view(E, N)[lo:hi], Q.view(E, N)[lo:hi], lo, hi, num_fix,
                           group)
    return [s.solve(max_iter, delta_thresh)]
