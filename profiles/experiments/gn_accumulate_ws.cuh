// Experimental warp-specialised persistent variant of the rays accumulate kernel (v8).
// NOT built by default (FRODO_RAYS_WS=0): on the 64-keyframe bench graph it is correct (same
// tests green) but slower than v7 -- 538 us vs 482 us per launch.  Ablations of this kernel
// (profiles/r1_ncu_accumulate.md): streams only 282 us, streams + arithmetic 371 us, streams +
// gathers 492 us: the gather path issued from 4 (or 8) producer warps is the limiter.
// Included by gn_accumulate.cu inside its anonymous namespace when FRODO_RAYS_WS=1.
#pragma once

// ---------------------------------------------------------------------------
// Warp-specialised persistent rays kernel (v8).
// One CTA per SM: 16 consumer warps (512 threads, one point pair each per 1024-point tile) and
// one producer warpgroup.  The producer owns ALL data movement, so memory requests are issued
// at the rate buffers free up, independent of where the consumers are in their arithmetic
// (v7 issued a tile's gathers from the consumer warps, between two tiles of arithmetic: the
// memory-only and arithmetic-only times of that kernel, 377 and 306 us, added up to 482).
// Two rings in shared memory:
//   * index ring (kWsIdxStages x 9 KB): idx + valid of a tile, bulk-copied (UBLKCP) a few tiles
//     ahead; read only by the producer lanes, which turn them into gather offsets;
//   * tile ring (kWsRing x 37 KB): everything the consumers read of a tile -- XP of frame j, Q,
//     valid (bulk copies, transaction bytes) and the gathered frame-i points (16-byte LDGSTS
//     from all 128 producer lanes, each lane arriving through cp.async.mbarrier.arrive.noinc
//     once its copies have landed) -- completing ONE mbarrier per tile.
// Consumers: wait that barrier, read operands, release the tile, rays_pair.  A CTA walks its
// work items (edge, chunk) back to back, so the rings never drain between items; registers
// are moved from the producer warpgroup to the consumers (setmaxnreg).
// ---------------------------------------------------------------------------
constexpr int kWsConsumers = 512;
constexpr int kWsProducers = 128;
constexpr int kWsThreads = kWsConsumers + kWsProducers;
constexpr int kWsTile = 2 * kWsConsumers;          // 1024 points
static_assert(!FRODO_RAYS_WS || kWsTile == kTilePts, "pack_points layout follows kTilePts");
constexpr int kWsIdxStages = 3;
constexpr int kWsIdxVm = kWsTile * 8;               // int64 [1024] then uint8 [1024]
constexpr int kWsIdxBytes = kWsIdxVm + kWsTile;
constexpr int kWsRing = 5;
#ifndef FRODO_WS_LAG
#define FRODO_WS_LAG 1
#endif
constexpr int kWsLag = 2;  // tiles of gathers a producer lane keeps in flight behind the newest
constexpr int kWsOffXP = 0;                         // float [512][8], pair-interleaved frame j
constexpr int kWsOffQ = kWsOffXP + kWsTile * 16;    // float [1024]
constexpr int kWsOffVm = kWsOffQ + kWsTile * 4;     // uint8 [1024]
constexpr int kWsOffGA = kWsOffVm + kWsTile;        // float4 [512] frame i, first points of the pairs
constexpr int kWsOffGB = kWsOffGA + kWsConsumers * 16;  // float4 [512] second points
constexpr int kWsTileBytes = kWsOffGB + kWsConsumers * 16;
static_assert(kWsTileBytes % 128 == 0 && kWsIdxBytes % 128 == 0, "ring alignment");
constexpr int kWsSmemBytes = kWsRing * kWsTileBytes + kWsIdxStages * kWsIdxBytes;
// launch: 640 threads x 96 registers; the pool is the CTA's own allocation:
// 512 x 112 + 128 x 32 = 640 x 96
constexpr int kWsConsumerRegs = 112, kWsProducerRegs = 32;

struct WsItem {
  int e, chunk, k_begin, npts, ntile;
};
__device__ __forceinline__ WsItem ws_item(int item, int nchunks, int pts_per_cta, int N) {
  WsItem w;
  w.e = item / nchunks;
  w.chunk = item - w.e * nchunks;
  w.k_begin = w.chunk * pts_per_cta;
  w.npts = min(N, w.k_begin + pts_per_cta) - w.k_begin;
  w.ntile = (w.npts + kWsTile - 1) / kWsTile;
  return w;
}

__global__ void __launch_bounds__(kWsThreads, 1)
gn_accumulate_rays_ws_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                             const float4* __restrict__ XC, const float4* __restrict__ XP,
                             const int32_t* __restrict__ ii_row,
                             const int32_t* __restrict__ jj_row, const int64_t* __restrict__ idx,
                             const uint8_t* __restrict__ valid_match, const float* __restrict__ Q,
                             int N, int pts_per_cta, int nchunks, int nitems,
                             float* __restrict__ partials, int32_t* __restrict__ tickets,
                             double* __restrict__ edge_blocks, const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  extern __shared__ __align__(128) unsigned char ring_raw[];
  __shared__ __align__(8) uint64_t ifull_bar[kWsIdxStages];
  __shared__ __align__(8) uint64_t iempty_bar[kWsIdxStages];
  __shared__ int cur[4];
  __shared__ __align__(8) uint64_t tfull_bar[kWsRing];
  __shared__ __align__(8) uint64_t tempty_bar[kWsRing];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < kWsIdxStages; ++s) {
      ptx::mbar_init(&ifull_bar[s], 1);
      ptx::mbar_init(&iempty_bar[s], kWsProducers / 32);
    }
    for (int g = 0; g < kWsRing; ++g) {
      ptx::mbar_init(&tfull_bar[g], kWsProducers + 1);  // gather lanes + the bulk-copy arrival
      ptx::mbar_init(&tempty_bar[g], kWsConsumers);
    }
    ptx::mbar_fence_init();
  }
  __syncthreads();
  unsigned char* idx_raw = ring_raw + kWsRing * kWsTileBytes;
  const uint32_t sm_ring = ptx::smem_u32(ring_raw);
  const uint32_t sm_ifull = ptx::smem_u32(&ifull_bar[0]);
  const uint32_t sm_iempty = ptx::smem_u32(&iempty_bar[0]);
  const uint32_t sm_tfull = ptx::smem_u32(&tfull_bar[0]);
  const uint32_t sm_tempty = ptx::smem_u32(&tempty_bar[0]);
  const int NB = (N + kWsTile - 1) / kWsTile;

  if (tid >= kWsConsumers) {
    // ------------------------------ producer warpgroup ------------------------------
    ptx::setmaxnreg_dec<kWsProducerRegs>();
    const int ptid = tid - kWsConsumers;
    const int pl = ptid & 31;
    // Cursor of the index stream, kWsIdxStages tiles ahead of the gathers.  Only lane 0 of
    // producer warp 0 uses it; it lives in shared memory so it costs no registers in the
    // gather lanes.
    auto issue_index = [&]() {  // idx + valid of the cursor's tile -> index stage; advance
      int c_item = cur[0], c_tile = cur[1], c_s = cur[2];
      if (c_item >= nitems) return;
      const WsItem cw = ws_item(c_item, nchunks, pts_per_cta, N);
      const int k0 = cw.k_begin + c_tile * kWsTile;
      const uint32_t np = (uint32_t)min(kWsTile, cw.npts - c_tile * kWsTile);
      unsigned char* st = idx_raw + (size_t)c_s * kWsIdxBytes;
      uint64_t* bar = &ifull_bar[c_s];
      const size_t eo = (size_t)cw.e * N + k0;
      ptx::mbar_arrive_expect_tx(bar, np * 9u);
      ptx::bulk_g2s(st, idx + eo, np * 8u, bar);
      ptx::bulk_g2s(st + kWsIdxVm, valid_match + eo, np, bar);
      c_s = c_s + 1 == kWsIdxStages ? 0 : c_s + 1;
      if (++c_tile == cw.ntile) {
        c_tile = 0;
        c_item += gridDim.x;
      }
      cur[0] = c_item; cur[1] = c_tile; cur[2] = c_s;
    };
    if (ptid == 0) {
      cur[0] = blockIdx.x; cur[1] = 0; cur[2] = 0;
      for (int s = 0; s < kWsIdxStages; ++s) issue_index();
    }

    uint32_t ifull_phase = 0, tempty_phase = 0, iempty_phase = 0;
    int si = 0, g = 0;  // index stage / ring slot of tile u
    int u = 0;
#if FRODO_WS_LAG
    int g_lag = 0;      // ring slot of tile u - kWsLag
#endif
    // Lane ptid gathers points k * 128 + ptid, k = 0..7 (0..511: first points of the pairs,
    // 512..1023: second points): consecutive lanes fetch the matches of consecutive points,
    // which are neighbours in frame i more often than not, so a copy instruction touches few
    // sectors (8 consecutive points per lane instead: 683 us instead of 540).
    constexpr int kPer = kWsTile / kWsProducers;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const WsItem w = ws_item(item, nchunks, pts_per_cta, N);
      const float4* __restrict__ XCi = XC + (size_t)ii_row[w.e] * N;
      for (int t = 0; t < w.ntile; ++t, ++u) {
        // 1. gather offsets from the index ring (landed tiles ago); the index stage goes back
        //    before the ring slot is awaited
        ptx::mbar_wait(sm_ifull + 8u * si, (ifull_phase >> si) & 1u);
        ifull_phase ^= 1u << si;
        const int left = w.npts - t * kWsTile;
        const unsigned char* st = idx_raw + si * kWsIdxBytes;
        int off[kPer];  // element offset into frame i (its low word: N < 2^31)
        {
          const int* ip = reinterpret_cast<const int*>(st) + 2 * ptid;
          const unsigned char* vp = st + kWsIdxVm + ptid;
          int lo[kPer];
          unsigned char vm[kPer];
#pragma unroll
          for (int k = 0; k < kPer; ++k) {
            lo[k] = ip[k * (2 * kWsProducers)];
            vm[k] = vp[k * kWsProducers];
          }
#pragma unroll
          for (int k = 0; k < kPer; ++k) off[k] = vm[k] ? lo[k] : 0;  // gn_kernels.cu:677
        }
        __syncwarp();
        if (pl == 0) ptx::mbar_arrive(sm_iempty + 8u * si);
        // 2. ring slot g: free once the consumers have read tile u - kWsRing
        if (u >= kWsRing) {
          ptx::mbar_wait(sm_tempty + 8u * g, (tempty_phase >> g) & 1u);
          tempty_phase ^= 1u << g;
        }
        const uint32_t slot = sm_ring + (uint32_t)g * kWsTileBytes;
        const uint32_t dst = slot + kWsOffGA + (uint32_t)ptid * 16u;
#pragma unroll
        for (int k = 0; k < kPer; ++k)
          if (k * kWsProducers + ptid < left && !(FRODO_ABL & 2))
            ptx::cp_async_16(dst + (uint32_t)k * (kWsProducers * 16u), XCi + off[k]);
#if FRODO_WS_LAG
        // completion of this lane's copies of tile u - kWsLag: by commit group, then a plain
        // arrival (the LDGSTS-tracking arrival waits for every outstanding copy of the lane)
        ptx::cp_async_commit();
        ptx::cp_async_wait<kWsLag>();
        if (u >= kWsLag) ptx::mbar_arrive(sm_tfull + 8u * g_lag);
        if (u >= kWsLag) g_lag = g_lag + 1 == kWsRing ? 0 : g_lag + 1;
#else
        ptx::cp_async_mbar_arrive_noinc(sm_tfull + 8u * g);
#endif
        if (ptid == 0) {
          const uint32_t np = (uint32_t)min(kWsTile, left);
          unsigned char* sl = ring_raw + (size_t)g * kWsTileBytes;
          uint64_t* tb = &tfull_bar[g];
          const size_t eo = (size_t)w.e * N + w.k_begin + (size_t)t * kWsTile;
          ptx::mbar_arrive_expect_tx(tb, kWsTile * 16u + np * 5u);
          ptx::bulk_g2s(sl + kWsOffXP,
                        XP + (size_t)jj_row[w.e] * NB * kWsTile + w.k_begin + (size_t)t * kWsTile,
                        kWsTile * 16u, tb);
          ptx::bulk_g2s(sl + kWsOffQ, Q + eo, np * 4u, tb);
          ptx::bulk_g2s(sl + kWsOffVm, valid_match + eo, np, tb);
          // all four producer warps have read index stage si: stream the next tile into it
          ptx::mbar_wait(sm_iempty + 8u * si, (iempty_phase >> si) & 1u);
          iempty_phase ^= 1u << si;
          issue_index();
        }
        si = si + 1 == kWsIdxStages ? 0 : si + 1;
        g = g + 1 == kWsRing ? 0 : g + 1;
      }
    }
    ptx::cp_async_wait<0>();  // no copies in flight at exit
#if FRODO_WS_LAG
    for (int k = u < kWsLag ? u : kWsLag; k > 0; --k) {
      ptx::mbar_arrive(sm_tfull + 8u * g_lag);
      g_lag = g_lag + 1 == kWsRing ? 0 : g_lag + 1;
    }
#endif
    return;
  }

  // -------------------------------- consumer warps --------------------------------
  ptx::setmaxnreg_inc<kWsConsumerRegs>();
  uint32_t tfull_phase = 0;
  int g = 0;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const WsItem w = ws_item(item, nchunks, pts_per_cta, N);
    const int ix = ii_row[w.e], jx = jj_row[w.e];
    const Sim3f Ti = load_pose(Twc, ix);
    RaysConst c;
    {
      const Sim3f Tj = load_pose(Twc, jx);
      const Sim3f Tij = rel_sim3(Ti, Tj);
      const float qx = Tij.q[0], qy = Tij.q[1], qz = Tij.q[2], qw = Tij.q[3], sc = Tij.s;
      // same map as actSO3 (no renormalisation of q), times the relative scale
      c.M[0] = sc * (1.f - 2.f * (qy * qy + qz * qz));
      c.M[1] = sc * (2.f * (qx * qy - qz * qw));
      c.M[2] = sc * (2.f * (qx * qz + qy * qw));
      c.M[3] = sc * (2.f * (qx * qy + qz * qw));
      c.M[4] = sc * (1.f - 2.f * (qx * qx + qz * qz));
      c.M[5] = sc * (2.f * (qy * qz - qx * qw));
      c.M[6] = sc * (2.f * (qx * qz - qy * qw));
      c.M[7] = sc * (2.f * (qy * qz + qx * qw));
      c.M[8] = sc * (1.f - 2.f * (qx * qx + qy * qy));
      c.t[0] = Tij.t[0]; c.t[1] = Tij.t[1]; c.t[2] = Tij.t[2];
      const float sa = 1.0f / prm.sigma_a, sb = 1.0f / prm.sigma_b;
      c.sa2 = sa * sa;
      c.sb2 = sb * sb;
      c.C_thresh = prm.C_thresh;
      c.Q_thresh = prm.Q_thresh;
    }
    float2 S2[28], v2[7];
#pragma unroll
    for (int k = 0; k < 28; ++k) S2[k] = f2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 7; ++k) v2[k] = f2(0.f, 0.f);
    int left = w.npts;
    for (int t = 0; t < w.ntile; ++t) {
      ptx::mbar_wait(sm_tfull + 8u * g, (tfull_phase >> g) & 1u);
      tfull_phase ^= 1u << g;
      if (tid < left) {
        const unsigned char* sl = ring_raw + g * kWsTileBytes;
        float4 ja = reinterpret_cast<const float4*>(sl + kWsOffXP)[2 * tid];
        float4 jb = reinterpret_cast<const float4*>(sl + kWsOffXP)[2 * tid + 1];
        const float* qp = reinterpret_cast<const float*>(sl + kWsOffQ);
        const float4 xiA = reinterpret_cast<const float4*>(sl + kWsOffGA)[tid];
        float4 xiB;
        float2 q;
        bool vmA = sl[kWsOffVm + tid] != 0, vmB;
        if (tid + kWsConsumers < left) {
          xiB = reinterpret_cast<const float4*>(sl + kWsOffGB)[tid];
          q = f2(qp[tid], qp[tid + kWsConsumers]);
          vmB = sl[kWsOffVm + tid + kWsConsumers] != 0;
        } else {  // ragged tail: lane y repeats lane x with zero weight
          xiB = xiA;
          q = f2(qp[tid], qp[tid]);
          vmB = false;
          ja = make_float4(ja.x, ja.x, ja.z, ja.z);
          jb = make_float4(jb.x, jb.x, jb.z, jb.z);
        }
        ptx::mbar_arrive(sm_tempty + 8u * g);  // operands are in registers
#if FRODO_ABL & 1
        S2[0] = fma2(f2(xiA.x, xiB.y), f2(ja.x, jb.y), S2[0]);
        S2[1] = fma2(q, f2(vmA ? 1.f : 0.f, vmB ? 1.f : 0.f), S2[1]);
#else
        rays_pair(c, S2, v2, xiA, xiB, f2(ja.x, ja.y), f2(ja.z, ja.w), f2(jb.x, jb.y),
                  f2(jb.z, jb.w), q, vmA, vmB);
#endif
      } else {
        ptx::mbar_arrive(sm_tempty + 8u * g);
      }
      g = g + 1 == kWsRing ? 0 : g + 1;
      left -= kWsTile;
    }
    float S[28], v[7];
#pragma unroll
    for (int k = 0; k < 28; ++k) S[k] = S2[k].x + S2[k].y;
#pragma unroll
    for (int k = 0; k < 7; ++k) v[k] = v2[k].x + v2[k].y;
    edge_epilogue<kWsConsumers, 1>(S, v, Ti, w.e, w.chunk, nchunks, nchunks, partials, tickets,
                                   edge_blocks);
  }
}

