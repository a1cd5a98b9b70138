// refine_matches: coarse-to-fine local descriptor search.
// Restates matching_kernels.cu:25-85.  For scalar_t = c10::Half the reference
// computes float(a)*float(b) -> fp16 (RN) and float(s)+float(p) -> fp16 (RN)
// for every k in ascending order; the exact product / sum of two fp16 values
// rounded once (HMUL2 / HADD2) gives the same bits (the fp32 intermediate is
// exact for the product and innocuous for the sum, 24 >= 2*11+2).  Candidates
// are visited in the reference order (dilation high->low, u outer, v inner,
// strict '>' keeps the first maximum, max_score starts at 0 for fp16 and at
// FLT_MIN for fp32, :48).
#include <limits.h>

#include <stdlib.h>

#include "common.cuh"
#include "ptx_async.cuh"

namespace frodo {

namespace {

// Two candidates are scored at once: lane halves (lo, hi) = (candidate c, c+1).
template <int F>
__device__ __forceinline__ __half2 score_pair_f16(const __half2* __restrict__ q2,  // [F/2] (q_2m, q_2m+1)
                                                  const __half* __restrict__ rowA,
                                                  const __half* __restrict__ rowB) {
  static_assert(F % 8 == 0, "vector path needs F % 8 == 0");
  __half2 s = __float2half2_rn(0.0f);
#pragma unroll
  for (int blk = 0; blk < F / 8; ++blk) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(rowA) + blk);
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(rowB) + blk);
    const unsigned aw[4] = {a.x, a.y, a.z, a.w};
    const unsigned bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      const __half2 av = *reinterpret_cast<const __half2*>(&aw[m]);
      const __half2 bv = *reinterpret_cast<const __half2*>(&bw[m]);
      const __half2 qv = q2[blk * 4 + m];
      const __half2 lo = __lows2half2(av, bv);    // (a_2m,   b_2m)
      const __half2 hi = __highs2half2(av, bv);   // (a_2m+1, b_2m+1)
      s = __hadd2_rn(s, __hmul2_rn(__low2half2(qv), lo));
      s = __hadd2_rn(s, __hmul2_rn(__high2half2(qv), hi));
    }
  }
  return s;
}

template <int F>
__global__ void __launch_bounds__(128)
refine_f16_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                  const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new,
                  __half* __restrict__ scores, int H, int W, int N, int radius, int dil_max) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const size_t b = blockIdx.y;
  const size_t t = b * (size_t)N + n;
  const __half* __restrict__ img = D11 + b * (size_t)H * W * F;

  __half2 q2[F / 2];
  {
    const uint4* qp = reinterpret_cast<const uint4*>(D21 + t * F);
#pragma unroll
    for (int blk = 0; blk < F / 8; ++blk) {
      const uint4 x = __ldg(qp + blk);
      q2[blk * 4 + 0] = *reinterpret_cast<const __half2*>(&x.x);
      q2[blk * 4 + 1] = *reinterpret_cast<const __half2*>(&x.y);
      q2[blk * 4 + 2] = *reinterpret_cast<const __half2*>(&x.z);
      q2[blk * 4 + 3] = *reinterpret_cast<const __half2*>(&x.w);
    }
  }
  const longlong2 p = __ldg(reinterpret_cast<const longlong2*>(p1) + t);
  long long u0 = p.x, v0 = p.y;
  long long u_new = u0, v_new = v0;
  __half max_score = __float2half_rn(0.0f);
  const int side = 2 * radius + 1;
  const int ncand = side * side;

  for (int d = dil_max; d > 0; --d) {
    const int rd = radius * d;
    for (int c = 0; c < ncand; c += 2) {
      // candidate c -> (i = c / side, j = c % side); u = u0 - rd + i*d, v = v0 - rd + j*d
      const int iA = c / side, jA = c - iA * side;
      const int cB = c + 1;
      const int iB = cB / side, jB = cB - iB * side;
      const long long uA = u0 - rd + (long long)iA * d, vA = v0 - rd + (long long)jA * d;
      const long long uB = u0 - rd + (long long)iB * d, vB = v0 - rd + (long long)jB * d;
      const bool inA = vA >= 0 && vA < H && uA >= 0 && uA < W;
      const bool inB = cB < ncand && vB >= 0 && vB < H && uB >= 0 && uB < W;
      if (!inA && !inB) continue;
      const __half* rowA = img + ((size_t)(inA ? vA : vB) * W + (size_t)(inA ? uA : uB)) * F;
      const __half* rowB = img + ((size_t)(inB ? vB : vA) * W + (size_t)(inB ? uB : uA)) * F;
      const __half2 s = score_pair_f16<F>(q2, rowA, rowB);
      const __half sA = __low2half(s), sB = __high2half(s);
      if (inA && __hgt(sA, max_score)) {
        max_score = sA;
        u_new = uA;
        v_new = vA;
      }
      if (inB && __hgt(sB, max_score)) {
        max_score = sB;
        u_new = uB;
        v_new = vB;
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  reinterpret_cast<longlong2*>(p1_new)[t] = make_longlong2(u_new, v_new);
  scores[t] = max_score;
}

// ---------------------------------------------------------------------------------------
// Tile-staged variant (the hot one): a CTA owns a 32x16 patch of query pixels, finds the
// bounding box of their search centres, and stages that box plus the level-dil_max halo of
// D11 in shared memory with one cp.async.bulk per tile row (the AoS rows of a box are
// contiguous in global memory).  A candidate row is then one conflict-free LDS.128 per 16
// bytes (lane stride = one pixel = 6F/8 banks x 4, which tiles the 32 banks per quarter
// warp for F = 24) instead of a line-granular L1 gather; candidates that fall outside the
// staged box (incoherent centres, box clipped to the shared-memory capacity) read global
// memory, so the result never depends on the staging.  Scores, visiting order and
// comparisons are those of refine_f16_kernel above.
constexpr int kTilePatchW = 32, kTilePatchH = 32, kTileThreads = kTilePatchW * kTilePatchH;
constexpr int kTileSmemBytes = 225 * 1024;
// Default for F <= 24: 16x16 patches.  Their box is 46x46 pixels (101 KB at F = 24), so TWO CTAs
// share an SM: one computes while the other stages its box (the 32x32 kernel, one CTA per SM,
// idles during every load), and a small grid -- one 512x384 pair is 192 CTAs of 32x32 on 148 SMs
// -- no longer leaves most of its last wave empty.  Measured, 512x384: 1 pair 212 -> 139 us,
// 8 pairs 852 -> 755 us, 256 pairs 2.11 -> 2.22 G px/s; bit-identical (tested against each other).
// 32x32 patches remain for F = 32 (the small box would not hold the halo).
constexpr int kSmallPatch = 16;
constexpr int kSmallSmemBytes = 112 * 1024;

template <int F>
__device__ __forceinline__ void load_row_f16(uint4 (&r)[F / 8], const unsigned char* tile,
                                             const __half* __restrict__ img, int u, int v,
                                             int tu0, int tv0, int TW, int TH, int pitch,
                                             int W) {
  const int ru = u - tu0, rv = v - tv0;
  if ((unsigned)ru < (unsigned)TW && (unsigned)rv < (unsigned)TH) {
    const uint4* p = reinterpret_cast<const uint4*>(tile + (size_t)(rv * pitch + ru) * (F * 2));
#pragma unroll
    for (int blk = 0; blk < F / 8; ++blk) r[blk] = p[blk];
  } else {
    const uint4* p = reinterpret_cast<const uint4*>(img + ((size_t)v * W + (size_t)u) * F);
#pragma unroll
    for (int blk = 0; blk < F / 8; ++blk) r[blk] = __ldg(p + blk);
  }
}

// Scores two candidate rows held in registers; lane halves (lo, hi) = (row a, row b).
template <int F>
__device__ __forceinline__ __half2 score_rows_f16(const __half2 (&q2)[F / 2],
                                                  const uint4 (&a)[F / 8],
                                                  const uint4 (&b)[F / 8]) {
  __half2 s = __float2half2_rn(0.0f);
#pragma unroll
  for (int blk = 0; blk < F / 8; ++blk) {
    const unsigned aw[4] = {a[blk].x, a[blk].y, a[blk].z, a[blk].w};
    const unsigned bw[4] = {b[blk].x, b[blk].y, b[blk].z, b[blk].w};
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      const __half2 av = *reinterpret_cast<const __half2*>(&aw[m]);
      const __half2 bv = *reinterpret_cast<const __half2*>(&bw[m]);
      const __half2 qv = q2[blk * 4 + m];
      s = __hadd2_rn(s, __hmul2_rn(__low2half2(qv), __lows2half2(av, bv)));
      s = __hadd2_rn(s, __hmul2_rn(__high2half2(qv), __highs2half2(av, bv)));
    }
  }
  return s;
}

// One search level whose whole (2R+1)^2 lattice lies inside the staged tile (the common,
// interior case): no bounds tests, candidate addresses are compile-time multiples of the two
// lattice strides.  `best` receives the index of the new maximum (untouched if none).
template <int F, int R>
__device__ __forceinline__ void level_in_tile(const __half2 (&q2)[F / 2],
                                              const unsigned char* base, int step_u, int step_v,
                                              __half& max_score, int& best) {
  constexpr int S = 2 * R + 1, NC = S * S;
#pragma unroll
  for (int c = 0; c < NC; c += 2) {
    constexpr int kLast = NC - 1;
    const int cb = c + 1 < NC ? c + 1 : kLast;
    const uint4* pa = reinterpret_cast<const uint4*>(base + (c / S) * step_u + (c % S) * step_v);
    const uint4* pb = reinterpret_cast<const uint4*>(base + (cb / S) * step_u + (cb % S) * step_v);
    uint4 ra[F / 8], rb[F / 8];
#pragma unroll
    for (int blk = 0; blk < F / 8; ++blk) {
      ra[blk] = pa[blk];
      rb[blk] = pb[blk];
    }
    const __half2 sc = score_rows_f16<F>(q2, ra, rb);
    if (__hgt(__low2half(sc), max_score)) {
      max_score = __low2half(sc);
      best = c;
    }
    if (c + 1 < NC && __hgt(__high2half(sc), max_score)) {
      max_score = __high2half(sc);
      best = c + 1;
    }
  }
}

// R > 0: radius known at compile time (interior levels take level_in_tile); R == 0: any radius.
template <int F, int R, int PW = kTilePatchW, int PH = kTilePatchH>
__global__ void __launch_bounds__(PW * PH, (PW * PH >= 1024) ? 1 : 2)
refine_f16_tile_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                       const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new,
                       __half* __restrict__ scores, int H, int W, int radius, int dil_max,
                       int cap_px, const float* __restrict__ D21_f32, const RefineTail tail) {
  extern __shared__ __align__(128) unsigned char tile[];
  __shared__ uint64_t bar;
  __shared__ int box[4];  // min u, max u, min v, max v of the centres that can reach the image
  const int tid = threadIdx.x;
  const int x = blockIdx.x * PW + (tid & (PW - 1));
  const int y = blockIdx.y * PH + (tid / PW);
  const size_t b = blockIdx.z;
  const bool active = x < W && y < H;
  const size_t t = b * (size_t)H * W + (size_t)y * W + x;
  const __half* __restrict__ img = D11 + b * (size_t)H * W * F;
  const int halo = radius * dil_max;

  if (tid == 0) {
    box[0] = INT_MAX; box[1] = INT_MIN; box[2] = INT_MAX; box[3] = INT_MIN;
    ptx::mbar_init(&bar, 1);
    ptx::mbar_fence_init();
  }
  longlong2 p = make_longlong2(0, 0);
  if (active) p = __ldg(reinterpret_cast<const longlong2*>(p1) + t);
  // A centre further than the first-level reach from the image never sees an in-bounds
  // candidate (the centre only moves onto in-bounds candidates), so it keeps p and score 0.
  const bool reach = active && p.x >= -(long long)halo && p.x < (long long)W + halo &&
                     p.y >= -(long long)halo && p.y < (long long)H + halo;
  int u0 = reach ? (int)p.x : 0, v0 = reach ? (int)p.y : 0;
  __syncthreads();
  {
    int mnu = reach ? u0 : INT_MAX, mxu = reach ? u0 : INT_MIN;
    int mnv = reach ? v0 : INT_MAX, mxv = reach ? v0 : INT_MIN;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mnu = min(mnu, __shfl_xor_sync(0xffffffffu, mnu, o));
      mxu = max(mxu, __shfl_xor_sync(0xffffffffu, mxu, o));
      mnv = min(mnv, __shfl_xor_sync(0xffffffffu, mnv, o));
      mxv = max(mxv, __shfl_xor_sync(0xffffffffu, mxv, o));
    }
    if ((tid & 31) == 0 && mnu <= mxu) {
      atomicMin(&box[0], mnu); atomicMax(&box[1], mxu);
      atomicMin(&box[2], mnv); atomicMax(&box[3], mxv);
    }
  }
  __syncthreads();
  // Row pitch of the staged box: a multiple of 8 pixels, so that pixels one row apart share a
  // bank group only at equal u -- lanes of a quarter warp (consecutive u, rows differing by at
  // most one for near-identity warps) then never collide (simulated 1.01 vs 1.12 wavefronts).
  int tu0 = 0, tv0 = 0, TW = 0, TH = 0, pitch = 8;
  if (box[0] <= box[1]) {
    tu0 = max(box[0] - halo, 0);
    tv0 = max(box[2] - halo, 0);
    TW = min(box[1] + halo, W - 1) - tu0 + 1;
    TH = min(box[3] + halo, H - 1) - tv0 + 1;
    if (TW > (cap_px & ~7)) TW = cap_px & ~7;  // pathological spread: the rest reads global memory
    pitch = (TW + 7) & ~7;
    if (pitch * TH > cap_px) {
      const int keep = cap_px / pitch;
      tv0 += (TH - keep) / 2;
      TH = keep;
    }
  }
  if (tid < 32 && TH > 0) {
    const uint32_t row_bytes = (uint32_t)TW * (F * 2);
    if (tid == 0) ptx::mbar_arrive_expect_tx(&bar, row_bytes * (uint32_t)TH);
    __syncwarp();
    for (int r = tid; r < TH; r += 32)
      ptx::bulk_g2s(tile + (size_t)r * pitch * (F * 2), img + ((size_t)(tv0 + r) * W + tu0) * F,
                    row_bytes, &bar);
  }

  __half2 q2[F / 2];
  if (reach && D21_f32 != nullptr) {
    // fp32 query descriptors (frodo_match): D21.half() of matching.py:80 done here, round to
    // nearest even per element as the cast kernel does, so the separate pass over D21 disappears
    const float4* qp = reinterpret_cast<const float4*>(D21_f32 + t * F);
#pragma unroll
    for (int m = 0; m < F / 4; ++m) {
      const float4 xq = __ldg(qp + m);
      q2[2 * m] = __floats2half2_rn(xq.x, xq.y);
      q2[2 * m + 1] = __floats2half2_rn(xq.z, xq.w);
    }
  } else if (reach) {
    const uint4* qp = reinterpret_cast<const uint4*>(D21 + t * F);
#pragma unroll
    for (int blk = 0; blk < F / 8; ++blk) {
      const uint4 xq = __ldg(qp + blk);
      q2[blk * 4 + 0] = *reinterpret_cast<const __half2*>(&xq.x);
      q2[blk * 4 + 1] = *reinterpret_cast<const __half2*>(&xq.y);
      q2[blk * 4 + 2] = *reinterpret_cast<const __half2*>(&xq.z);
      q2[blk * 4 + 3] = *reinterpret_cast<const __half2*>(&xq.w);
    }
  } else {
#pragma unroll
    for (int m = 0; m < F / 2; ++m) q2[m] = __float2half2_rn(0.0f);
  }
  if (TH > 0) ptx::mbar_wait(&bar, 0);

  __half max_score = __float2half_rn(0.0f);
  if (reach) {
    const int side = 2 * radius + 1;
    const int ncand = side * side;
    for (int d = dil_max; d > 0; --d) {
      const int ub = u0 - radius * d, vb = v0 - radius * d;
      const int span = 2 * radius * d;
      const int ru = ub - tu0, rv = vb - tv0;
      if (R > 0 && ru >= 0 && ru + span < TW && rv >= 0 && rv + span < TH) {
        int best = -1;
        level_in_tile<F, (R > 0 ? R : 1)>(q2, tile + (size_t)(rv * pitch + ru) * (F * 2),
                                          d * (F * 2), d * pitch * (F * 2), max_score, best);
        if (best >= 0) {
          constexpr int S = 2 * (R > 0 ? R : 1) + 1;
          u0 = ub + (best / S) * d;
          v0 = vb + (best % S) * d;
        }
        continue;
      }
      // border / out-of-tile level: per-candidate bounds tests, global-memory fallback
      int u_new = u0, v_new = v0;
      int i = 0, j = 0;  // candidate c = i * side + j  ->  (ub + i*d, vb + j*d)
      for (int c = 0; c < ncand; c += 2) {
        int cu[2], cv[2];
        bool in[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          cu[k] = ub + i * d;
          cv[k] = vb + j * d;
          in[k] = c + k < ncand && cv[k] >= 0 && cv[k] < H && cu[k] >= 0 && cu[k] < W;
          if (++j == side) { j = 0; ++i; }
        }
        if (!in[0] && !in[1]) continue;
        // an out-of-image candidate borrows the address of the valid one; its score is ignored
        uint4 ra[F / 8], rb[F / 8];
        load_row_f16<F>(ra, tile, img, in[0] ? cu[0] : cu[1], in[0] ? cv[0] : cv[1], tu0, tv0, TW,
                        TH, pitch, W);
        load_row_f16<F>(rb, tile, img, in[1] ? cu[1] : cu[0], in[1] ? cv[1] : cv[0], tu0, tv0, TW,
                        TH, pitch, W);
        const __half2 s0 = score_rows_f16<F>(q2, ra, rb);
        if (in[0] && __hgt(__low2half(s0), max_score)) {
          max_score = __low2half(s0);
          u_new = cu[0];
          v_new = cv[0];
        }
        if (in[1] && __hgt(__high2half(s0), max_score)) {
          max_score = __high2half(s0);
          u_new = cu[1];
          v_new = cv[1];
        }
      }
      u0 = u_new;
      v0 = v_new;
    }
  }
  if (active) {
    const longlong2 best = reach ? make_longlong2((long long)u0, (long long)v0) : p;
    if (tail.valid_in == nullptr) {
      reinterpret_cast<longlong2*>(p1_new)[t] = best;
      scores[t] = max_score;
    } else {
      // frodo_match: matching.py:87-92 from the registers (as match_finish_kernel): the score is
      // compared in fp16 against the threshold cast to fp16
      tail.idx[t] = best.x + (long long)W * best.y;
      tail.valid[t] = (tail.valid_in[t] != 0) && __hgt(max_score, __float2half_rn(tail.threshold));
    }
  }
}

// Generic-F fp16 path (scalar, any F).
__global__ void __launch_bounds__(128)
refine_f16_generic_kernel(const __half* __restrict__ D11, const __half* __restrict__ D21,
                          const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new,
                          __half* __restrict__ scores, int H, int W, int N, int F, int radius,
                          int dil_max) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const size_t b = blockIdx.y;
  const size_t t = b * (size_t)N + n;
  const __half* __restrict__ img = D11 + b * (size_t)H * W * F;
  const __half* __restrict__ q = D21 + t * F;
  long long u0 = p1[t * 2], v0 = p1[t * 2 + 1];
  long long u_new = u0, v_new = v0;
  __half max_score = __float2half_rn(0.0f);
  for (int d = dil_max; d > 0; --d) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i, v = v0 - rd + j;
        if (v >= 0 && v < H && u >= 0 && u < W) {
          const __half* row = img + ((size_t)v * W + (size_t)u) * F;
          __half s = __float2half_rn(0.0f);
          for (int k = 0; k < F; ++k) s = __hadd_rn(s, __hmul_rn(q[k], row[k]));
          if (__hgt(s, max_score)) {
            max_score = s;
            u_new = u;
            v_new = v;
          }
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[t * 2] = u_new;
  p1_new[t * 2 + 1] = v_new;
  scores[t] = max_score;
}

// fp32 / fp64 path (the reference dispatches AT_DISPATCH_FLOATING_TYPES_AND_HALF,
// matching_kernels.cu:112): score += a*b is contracted to fma(a, b, score) by nvcc in the
// reference; max_score starts at numeric_limits<scalar_t>::min().
template <typename T>
__device__ __forceinline__ T refine_min_score();
template <>
__device__ __forceinline__ float refine_min_score<float>() { return 1.17549435e-38f; }
template <>
__device__ __forceinline__ double refine_min_score<double>() { return 2.2250738585072014e-308; }

template <typename T>
__global__ void __launch_bounds__(128)
refine_f32_kernel(const T* __restrict__ D11, const T* __restrict__ D21,
                  const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new,
                  T* __restrict__ scores, int H, int W, int N, int F, int radius,
                  int dil_max) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const size_t b = blockIdx.y;
  const size_t t = b * (size_t)N + n;
  const T* __restrict__ img = D11 + b * (size_t)H * W * F;
  const T* __restrict__ q = D21 + t * F;
  long long u0 = p1[t * 2], v0 = p1[t * 2 + 1];
  long long u_new = u0, v_new = v0;
  T max_score = refine_min_score<T>();
  for (int d = dil_max; d > 0; --d) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i, v = v0 - rd + j;
        if (v >= 0 && v < H && u >= 0 && u < W) {
          const T* row = img + ((size_t)v * W + (size_t)u) * F;
          T s = (T)0;
          for (int k = 0; k < F; ++k) s = fma(__ldg(q + k), __ldg(row + k), s);
          if (s > max_score) {
            max_score = s;
            u_new = u;
            v_new = v;
          }
        }
      }
    }
    u0 = u_new;
    v0 = v_new;
  }
  p1_new[t * 2] = u_new;
  p1_new[t * 2 + 1] = v_new;
  scores[t] = max_score;
}

}  // namespace

// True when launch_refine_f16 takes the tile-staged kernel for these arguments: only that kernel
// can read fp32 query descriptors (D21_f32), so frodo_match asks before skipping the D21 cast.
bool refine_f16_takes_tile_kernel(const void* D11, const void* D21, const void* p1,
                                  const void* p1_new, int B, int H, int W, int N, int F) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21) |
                         reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(p1_new)) &
                        15) == 0;
  return aligned && N == H * W && (F == 24 || F == 16 || F == 32) && B <= 65535;
}

// Patch-size override of the tile-staged kernel: 0 = automatic, 16 or 32.  The environment
// variable FRODO_REFINE_PATCH seeds it once per process; frodo_set_refine_patch() changes it at
// run time (tests compare the two variants bit for bit).  set < 0 only reads.
int refine_forced_patch(int set) {
  static int value = [] {
    const char* f = getenv("FRODO_REFINE_PATCH");
    return f == nullptr ? 0 : (f[0] == '1' ? 16 : (f[0] == '3' ? 32 : 0));
  }();
  if (set >= 0) value = (set == 16 || set == 32) ? set : 0;
  return value;
}

int launch_refine_f16(const __half* D11, const __half* D21, const int64_t* p1, int64_t* p1_new,
                      __half* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s, const float* D21_f32, const RefineTail* tail_in) {
  if (B == 0 || N == 0) return FRODO_OK;
  dim3 grid((N + 127) / 128, B);
  const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21) |
                         reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(p1_new)) &
                        15) == 0;
  const bool tile = refine_f16_takes_tile_kernel(D11, D21_f32 ? (const void*)D21_f32 : (const void*)D21,
                                                 p1, p1_new, B, H, W, N, F);
  RefineTail tail = {};
  if (tail_in != nullptr) tail = *tail_in;
  if ((D21_f32 != nullptr || tail.valid_in != nullptr) && !tile) {
    set_error("refine_matches: fp32 query descriptors / the fused tail need the tile-staged kernel");
    return FRODO_ERR_ARG;
  }
  if (tile) {
    const long long tiles32 = (long long)((W + kTilePatchW - 1) / kTilePatchW) *
                              ((H + kTilePatchH - 1) / kTilePatchH) * B;
    (void)tiles32;
    bool small_grid = F <= 24;  // "small" = the 16x16, two-CTAs-per-SM variant
    const int forced_patch = refine_forced_patch(-1);  // tuning/test switch, see below
    if (forced_patch == 16 && F <= 24) small_grid = true;
    if (forced_patch == 32) small_grid = false;
    const int pw = small_grid ? kSmallPatch : kTilePatchW, ph = small_grid ? kSmallPatch : kTilePatchH;
    const int smem_bytes = small_grid ? kSmallSmemBytes : kTileSmemBytes;
    dim3 tgrid((W + pw - 1) / pw, (H + ph - 1) / ph, B);
    const int cap_px = smem_bytes / (F * 2);
#define FRODO_TILE_LAUNCH_P(FF, RR, PWW, PHH)                                                   \
  do {                                                                                          \
    static PerDeviceOnce once;                                                                  \
    int dev_;                                                                                   \
    if (once.needed(&dev_)) {                                                                   \
      FRODO_CUDA_TRY(cudaFuncSetAttribute(refine_f16_tile_kernel<FF, RR, PWW, PHH>,             \
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                          smem_bytes));                                         \
      once.mark(dev_);                                                                          \
    }                                                                                           \
    refine_f16_tile_kernel<FF, RR, PWW, PHH><<<tgrid, PWW * PHH, smem_bytes, s>>>(              \
        D11, D21, p1, p1_new, scores, H, W, radius, dil, cap_px, D21_f32, tail);                \
  } while (0)
#define FRODO_TILE_LAUNCH(FF, RR)                                                               \
  do {                                                                                          \
    if (small_grid) FRODO_TILE_LAUNCH_P(FF, RR, kSmallPatch, kSmallPatch);                      \
    else FRODO_TILE_LAUNCH_P(FF, RR, kTilePatchW, kTilePatchH);                                 \
  } while (0)
    if (F == 24 && radius == 3) FRODO_TILE_LAUNCH(24, 3);
    else if (F == 24) FRODO_TILE_LAUNCH(24, 0);
    else if (F == 16 && radius == 3) FRODO_TILE_LAUNCH(16, 3);
    else if (F == 16) FRODO_TILE_LAUNCH(16, 0);
    else if (radius == 3) FRODO_TILE_LAUNCH(32, 3);
    else FRODO_TILE_LAUNCH(32, 0);
#undef FRODO_TILE_LAUNCH_P
#undef FRODO_TILE_LAUNCH
  } else if (F == 24 && aligned) {
    refine_f16_kernel<24><<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, radius, dil);
  } else if (F == 16 && aligned) {
    refine_f16_kernel<16><<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, radius, dil);
  } else if (F == 32 && aligned) {
    refine_f16_kernel<32><<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, radius, dil);
  } else {
    refine_f16_generic_kernel<<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, F,
                                                   radius, dil);
  }
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_refine_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new,
                      float* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s) {
  if (B == 0 || N == 0) return FRODO_OK;
  dim3 grid((N + 127) / 128, B);
  refine_f32_kernel<float><<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, F, radius,
                                                dil);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_refine_f64(const double* D11, const double* D21, const int64_t* p1, int64_t* p1_new,
                      double* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s) {
  if (B == 0 || N == 0) return FRODO_OK;
  dim3 grid((N + 127) / 128, B);
  refine_f32_kernel<double><<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, F, radius,
                                                 dil);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
