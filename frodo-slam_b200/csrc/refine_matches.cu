// refine_matches: coarse-to-fine local descriptor search.
// Restates matching_kernels.cu:25-85.  For scalar_t = c10::Half the reference
// computes float(a)*float(b) -> fp16 (RN) and float(s)+float(p) -> fp16 (RN)
// for every k in ascending order; the exact product / sum of two fp16 values
// rounded once (HMUL2 / HADD2) gives the same bits (the fp32 intermediate is
// exact for the product and innocuous for the sum, 24 >= 2*11+2).  Candidates
// are visited in the reference order (dilation high->low, u outer, v inner,
// strict '>' keeps the first maximum, max_score starts at 0 for fp16 and at
// FLT_MIN for fp32, :48).
#include "common.cuh"

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

// fp32 path: score += a*b is contracted to fma(a, b, score) by nvcc in the reference.
__global__ void __launch_bounds__(128)
refine_f32_kernel(const float* __restrict__ D11, const float* __restrict__ D21,
                  const int64_t* __restrict__ p1, int64_t* __restrict__ p1_new,
                  float* __restrict__ scores, int H, int W, int N, int F, int radius,
                  int dil_max) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const size_t b = blockIdx.y;
  const size_t t = b * (size_t)N + n;
  const float* __restrict__ img = D11 + b * (size_t)H * W * F;
  const float* __restrict__ q = D21 + t * F;
  long long u0 = p1[t * 2], v0 = p1[t * 2 + 1];
  long long u_new = u0, v_new = v0;
  float max_score = 1.17549435e-38f;  // numeric_limits<float>::min()
  for (int d = dil_max; d > 0; --d) {
    const int rd = radius * d;
    const int diam = 2 * rd + 1;
    for (int i = 0; i < diam; i += d) {
      for (int j = 0; j < diam; j += d) {
        const long long u = u0 - rd + i, v = v0 - rd + j;
        if (v >= 0 && v < H && u >= 0 && u < W) {
          const float* row = img + ((size_t)v * W + (size_t)u) * F;
          float s = 0.0f;
          for (int k = 0; k < F; ++k) s = __fmaf_rn(__ldg(q + k), __ldg(row + k), s);
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

int launch_refine_f16(const __half* D11, const __half* D21, const int64_t* p1, int64_t* p1_new,
                      __half* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s) {
  if (B == 0 || N == 0) return FRODO_OK;
  dim3 grid((N + 127) / 128, B);
  const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21) |
                         reinterpret_cast<uintptr_t>(p1) | reinterpret_cast<uintptr_t>(p1_new)) &
                        15) == 0;
  if (F == 24 && aligned) {
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
  refine_f32_kernel<<<grid, 128, 0, s>>>(D11, D21, p1, p1_new, scores, H, W, N, F, radius, dil);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
