// The matching pipeline around iter_proj / refine_matches: what the reference
// does with ~20 PyTorch kernels per call in mast3r_slam/matching.py:25-110 and
// mast3r_slam/image.py:5-38, as four streaming kernels.
//
//   match_prep_kernel       matching.py:25-49  F.normalize(X11) -> rays, Scharr gradient with
//                           reflect padding (image.py:5-38), concat to [b,h,w,9];
//                           F.normalize(X21) -> pts3d_norm; idx_1_to_2_init -> p_init (x, y)
//   match_occlusion_kernel  matching.py:66-75  p1 = trunc(p), valid = converged &
//                           ||X11[p1] - X21|| < dist_thresh
//   convert_f32_f16_kernel  matching.py:79-80  D.half()
//   match_finish_kernel     matching.py:87-92  valid &= score > threshold, idx = x + w * y
//
// All four are HBM-bound element-wise passes (80 / 50 / 6 per element / 28 bytes per pixel).
// The operation order is fixed and documented in include/frodo_b200.h -- separately rounded
// products and sums (__fmul_rn/__fadd_rn so that nvcc does not contract them), IEEE sqrt and
// division -- so the outputs are reproducible bit for bit by the CPU restatement the tests
// check against; torch's kernels order the same sums differently (1-2 ulp).
#include "common.cuh"

namespace frodo {
namespace {

__device__ __forceinline__ int reflect(int i, int n) {  // F.pad(..., mode="reflect"), pad 1
  return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
}

// x / max(||x||, 1e-12) with n2 = (x0 x0 + x1 x1) + x2 x2
__device__ __forceinline__ float3 normalize3(float x, float y, float z) {
  const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
  const float d = fmaxf(__fsqrt_rn(n2), 1e-12f);
  return make_float3(__fdiv_rn(x, d), __fdiv_rn(y, d), __fdiv_rn(z, d));
}

constexpr int kPrepTW = 32, kPrepTH = 8;  // pixels per CTA: one per thread
constexpr int kPrepThreads = kPrepTW * kPrepTH;
constexpr int kHaloW = kPrepTW + 2, kHaloH = kPrepTH + 2;

__global__ void __launch_bounds__(kPrepThreads)
match_prep_kernel(const float* __restrict__ X11, const float* __restrict__ X21,
                  const int64_t* __restrict__ idx_init, float* __restrict__ rays_with_grad,
                  float* __restrict__ pts3d_norm, float* __restrict__ p_init, int H, int W) {
  __shared__ float ray[kHaloH][kHaloW][3];            // normalised rays of the tile + halo
  __shared__ __align__(16) float outbuf[kPrepTH][kPrepTW * 9];  // staged for coalesced stores
  const int b = blockIdx.z;
  const int x0 = blockIdx.x * kPrepTW, y0 = blockIdx.y * kPrepTH;
  const int tid = threadIdx.x;
  const size_t img = (size_t)b * H * W;
  for (int i = tid; i < kHaloH * kHaloW; i += kPrepThreads) {
    const int hy = i / kHaloW, hx = i - hy * kHaloW;
    // positions beyond the image (ragged tiles) are clamped; they feed no written pixel
    const int yy = reflect(min(y0 + hy - 1, H), H), xx = reflect(min(x0 + hx - 1, W), W);
    const float* p = X11 + (img + (size_t)yy * W + xx) * 3;
    const float3 r = normalize3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
    ray[hy][hx][0] = r.x;
    ray[hy][hx][1] = r.y;
    ray[hy][hx][2] = r.z;
  }
  __syncthreads();
  const int tx = tid % kPrepTW, ty = tid / kPrepTW;
  const int x = x0 + tx, y = y0 + ty;
  const bool inside = x < W && y < H;
  {
    const float w3 = 3.0f / 32.0f, w10 = 10.0f / 32.0f;
    float* o = &outbuf[ty][tx * 9];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float a00 = ray[ty][tx][c], a01 = ray[ty][tx + 1][c], a02 = ray[ty][tx + 2][c];
      const float a10 = ray[ty + 1][tx][c], a12 = ray[ty + 1][tx + 2][c];
      const float a20 = ray[ty + 2][tx][c], a21 = ray[ty + 2][tx + 1][c], a22 = ray[ty + 2][tx + 2][c];
      float gx = __fmul_rn(a00, -w3);  // 0 + t is t
      gx = __fadd_rn(gx, __fmul_rn(a02, w3));
      gx = __fadd_rn(gx, __fmul_rn(a10, -w10));
      gx = __fadd_rn(gx, __fmul_rn(a12, w10));
      gx = __fadd_rn(gx, __fmul_rn(a20, -w3));
      gx = __fadd_rn(gx, __fmul_rn(a22, w3));
      float gy = __fmul_rn(a00, -w3);
      gy = __fadd_rn(gy, __fmul_rn(a01, -w10));
      gy = __fadd_rn(gy, __fmul_rn(a02, -w3));
      gy = __fadd_rn(gy, __fmul_rn(a20, w3));
      gy = __fadd_rn(gy, __fmul_rn(a21, w10));
      gy = __fadd_rn(gy, __fmul_rn(a22, w3));
      o[c] = ray[ty + 1][tx + 1][c];
      o[3 + c] = gx;
      o[6 + c] = gy;
    }
  }
  if (inside) {
    const size_t n = img + (size_t)y * W + x;
    const float* q = X21 + n * 3;
    const float3 r = normalize3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
    float* o = pts3d_norm + n * 3;
    o[0] = r.x;
    o[1] = r.y;
    o[2] = r.z;
    const int64_t id = idx_init != nullptr ? __ldg(idx_init + n) : (int64_t)y * W + x;
    // lin_to_pixel (matching.py:18-22): python floor division / modulo, then .float()
    int64_t v = id / W, u = id - v * W;
    if (u < 0) {
      u += W;
      v -= 1;
    }
    reinterpret_cast<float2*>(p_init)[n] = make_float2((float)u, (float)v);
  }
  __syncthreads();
  // rows of the tile are contiguous in [b,h,w,9]: kPrepTW * 9 floats each
  const int cols = min(kPrepTW, W - x0) * 9;
  for (int r = 0; r < kPrepTH && y0 + r < H; ++r) {
    float* dst = rays_with_grad + (img + (size_t)(y0 + r) * W + x0) * 9;
    if ((cols & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
      for (int i = tid; i < cols / 4; i += kPrepThreads)
        reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(&outbuf[r][0])[i];
    } else {
      for (int i = tid; i < cols; i += kPrepThreads) dst[i] = outbuf[r][i];
    }
  }
}

__global__ void __launch_bounds__(256)
match_occlusion_kernel(const float* __restrict__ X11, const float* __restrict__ X21,
                       const float2* __restrict__ p_iter, const uint8_t* __restrict__ conv,
                       float dist_thresh, longlong2* __restrict__ p1, uint8_t* __restrict__ valid,
                       int H, int W, size_t total) {
  const size_t hw = (size_t)H * W;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x; n < total; n += stride) {
    const size_t b = n / hw;
    const float2 p = __ldg(p_iter + n);
    const long long px = (long long)p.x, py = (long long)p.y;  // .long(): toward zero
    p1[n] = make_longlong2(px, py);
    // iter_proj keeps p inside the image; the clamp only guards the address
    const long long gx = min(max(px, 0LL), (long long)W - 1), gy = min(max(py, 0LL), (long long)H - 1);
    const float* a = X11 + (b * hw + (size_t)gy * W + (size_t)gx) * 3;
    const float* q = X21 + n * 3;
    const float d0 = __fsub_rn(__ldg(a), __ldg(q));
    const float d1 = __fsub_rn(__ldg(a + 1), __ldg(q + 1));
    const float d2 = __fsub_rn(__ldg(a + 2), __ldg(q + 2));
    const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2));
    valid[n] = (conv[n] != 0) && (__fsqrt_rn(n2) < dist_thresh);
  }
}

// 8 elements per thread and trip: two float4 in, one uint4 (8 halves) out; round to nearest even
__global__ void __launch_bounds__(256)
convert_f32_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t n8 = n / 8;
  const bool vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  if (vec) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
      const float4 a = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i);
      const float4 c = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i + 1);
      const __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w);
      const __half2 h2 = __floats2half2_rn(c.x, c.y), h3 = __floats2half2_rn(c.z, c.w);
      uint4 o;
      o.x = *reinterpret_cast<const uint32_t*>(&h0);
      o.y = *reinterpret_cast<const uint32_t*>(&h1);
      o.z = *reinterpret_cast<const uint32_t*>(&h2);
      o.w = *reinterpret_cast<const uint32_t*>(&h3);
      reinterpret_cast<uint4*>(dst)[i] = o;
    }
  }
  const size_t tail = vec ? n8 * 8 : 0;
  for (size_t i = tail + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2half_rn(src[i]);
}

template <typename ScoreT>
__global__ void __launch_bounds__(256)
match_finish_kernel(const longlong2* __restrict__ p1, const ScoreT* __restrict__ scores,
                    const uint8_t* __restrict__ valid_in, float threshold,
                    int64_t* __restrict__ idx, uint8_t* __restrict__ valid, int W, size_t total) {
  // `scores > threshold` is evaluated in the scores' dtype (the scalar is cast to it)
  const ScoreT thr = static_cast<ScoreT>(threshold);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x; n < total; n += stride) {
    const longlong2 p = p1[n];
    idx[n] = p.x + (long long)W * p.y;
    valid[n] = (valid_in[n] != 0) && (scores[n] > thr);
  }
}

// The per-edge quality glue the callers run after match(): tracker.py:194-206 and
// FactorGraph.add_factors, global_opt.py:98-112 (six torch kernels + two boolean sums there).
//   Q[n]         = sqrt(Q_src[b, idx[n]] * Q_dst[n])                    (fp32, one rounding each)
//   valid_q[n]   = valid_match[n] & (Q[n] > Q_conf)
//   valid_opt[n] = valid_q[n] & (C_src[b, idx[n]] > C_conf) & (C_dst[n] > C_conf)   (if C given)
//   counts[b]    = (sum valid_opt, sum valid_q): numerators of match_frac
// One thread per 4 consecutive points (16-byte streams); the two gathers are the only
// data-dependent accesses.  Algorithmic bytes per point: idx 8 + Q 4+4 + valid 1 (+ C 4+4) in,
// Q 4 + 2 masks out = 23 (31).
constexpr int kQualThreads = 256;
__global__ void __launch_bounds__(kQualThreads)
match_quality_kernel(const float* __restrict__ Q_src, const float* __restrict__ Q_dst,
                     const float* __restrict__ C_src, const float* __restrict__ C_dst,
                     const int64_t* __restrict__ idx, const uint8_t* __restrict__ valid_match,
                     float Q_conf, float C_conf, float* __restrict__ Q_out,
                     uint8_t* __restrict__ valid_opt, uint8_t* __restrict__ valid_q,
                     int32_t* __restrict__ counts, int N, int N_src) {
  const int b = blockIdx.y;
  const size_t row = (size_t)b * N;
  const float* qs = Q_src + (size_t)b * N_src;
  const float* cs = C_src ? C_src + (size_t)b * N_src : nullptr;
  int n_opt = 0, n_q = 0;
  const int stride = gridDim.x * kQualThreads * 4;
  for (int n0 = (blockIdx.x * kQualThreads + threadIdx.x) * 4; n0 < N; n0 += stride) {
    long long m[4];
    float qd[4], cd[4], qg[4], cg[4];
    bool vm[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int n = min(n0 + u, N - 1);  // clamped: loads stay unconditional
      m[u] = idx[row + n];
      qd[u] = Q_dst[row + n];
      vm[u] = valid_match[row + n] != 0;
      cd[u] = C_dst ? C_dst[row + n] : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      qg[u] = qs[m[u]];
      cg[u] = cs ? cs[m[u]] : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int n = n0 + u;
      if (n >= N) break;
      const float q = __fsqrt_rn(__fmul_rn(qg[u], qd[u]));
      const bool vq = vm[u] && (q > Q_conf);
      bool vo = vq;
      if (C_src) vo = vo && (cg[u] > C_conf);
      if (C_dst) vo = vo && (cd[u] > C_conf);
      Q_out[row + n] = q;
      valid_q[row + n] = vq;
      valid_opt[row + n] = vo;
      n_q += vq;
      n_opt += vo;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_opt += __shfl_xor_sync(0xffffffffu, n_opt, o);
    n_q += __shfl_xor_sync(0xffffffffu, n_q, o);
  }
  if ((threadIdx.x & 31) == 0 && (n_opt | n_q)) {
    atomicAdd(counts + 2 * b, n_opt);
    atomicAdd(counts + 2 * b + 1, n_q);
  }
}

int stream_blocks(size_t work_items) {
  const size_t want = (work_items + 255) / 256;
  const size_t cap = (size_t)num_sms() * 8;  // 8 resident CTAs of 256 threads per SM
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace

int launch_match_prep(const float* X11, const float* X21, const int64_t* idx_init,
                      float* rays_with_grad, float* pts3d_norm, float* p_init, int B, int H, int W,
                      cudaStream_t s) {
  if (B == 0) return FRODO_OK;
  dim3 grid((W + kPrepTW - 1) / kPrepTW, (H + kPrepTH - 1) / kPrepTH, B);
  match_prep_kernel<<<grid, kPrepThreads, 0, s>>>(X11, X21, idx_init, rays_with_grad, pts3d_norm,
                                                  p_init, H, W);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_match_occlusion(const float* X11, const float* X21, const float* p_iter,
                           const uint8_t* conv, float dist_thresh, int64_t* p1, uint8_t* valid,
                           int B, int H, int W, cudaStream_t s) {
  const size_t total = (size_t)B * H * W;
  if (total == 0) return FRODO_OK;
  match_occlusion_kernel<<<stream_blocks(total), 256, 0, s>>>(
      X11, X21, reinterpret_cast<const float2*>(p_iter), conv, dist_thresh,
      reinterpret_cast<longlong2*>(p1), valid, H, W, total);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

// constrain_points_to_ray (geometry.py:37-42 + backproject :107-115), what solve_GN_calib does
// to the pointmaps before gauss_newton_calib (global_opt.py:296-300): point n of a frame keeps
// its depth z and moves onto the ray of its pixel, X = z * ((u - cx) / fx, (v - cy) / fy, 1),
// u = n mod W, v = n / W.  IEEE subtract, divide, multiply: bit-identical to the torch ops.
namespace {
__global__ void __launch_bounds__(256)
constrain_points_kernel(const float* __restrict__ X, const float* __restrict__ Kmat, int HW,
                        int W, size_t total, float* __restrict__ out) {
  const float fx = __ldg(Kmat + 0), fy = __ldg(Kmat + 4), cx = __ldg(Kmat + 2), cy = __ldg(Kmat + 5);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x; n < total; n += stride) {
    const int pix = (int)(n % (size_t)HW);
    const float u = (float)(pix % W), v = (float)(pix / W);
    const float z = __ldg(X + n * 3 + 2);
    out[n * 3 + 0] = __fmul_rn(z, __fdiv_rn(__fsub_rn(u, cx), fx));
    out[n * 3 + 1] = __fmul_rn(z, __fdiv_rn(__fsub_rn(v, cy), fy));
    out[n * 3 + 2] = __fmul_rn(z, 1.0f);
  }
}
}  // namespace

int launch_constrain_points(const float* X, const float* Kmat, int num_frames, int H, int W,
                            float* out, cudaStream_t s) {
  const size_t total = (size_t)num_frames * H * W;
  if (total == 0) return FRODO_OK;
  constrain_points_kernel<<<stream_blocks(total), 256, 0, s>>>(X, Kmat, H * W, W, total, out);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_convert_f32_f16(const float* src, __half* dst, size_t n, cudaStream_t s) {
  if (n == 0) return FRODO_OK;
  convert_f32_f16_kernel<<<stream_blocks(n / 8 + 1), 256, 0, s>>>(src, dst, n);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_match_finish(const int64_t* p1, const void* scores, bool scores_f16,
                        const uint8_t* valid_in, float threshold, int64_t* idx, uint8_t* valid,
                        int B, int N, int W, cudaStream_t s) {
  const size_t total = (size_t)B * N;
  if (total == 0) return FRODO_OK;
  if (scores_f16)
    match_finish_kernel<__half><<<stream_blocks(total), 256, 0, s>>>(
        reinterpret_cast<const longlong2*>(p1), static_cast<const __half*>(scores), valid_in,
        threshold, idx, valid, W, total);
  else
    match_finish_kernel<float><<<stream_blocks(total), 256, 0, s>>>(
        reinterpret_cast<const longlong2*>(p1), static_cast<const float*>(scores), valid_in,
        threshold, idx, valid, W, total);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_match_quality(const float* Q_src, const float* Q_dst, const float* C_src,
                         const float* C_dst, const int64_t* idx, const uint8_t* valid_match,
                         float Q_conf, float C_conf, float* Q_out, uint8_t* valid_opt,
                         uint8_t* valid_q, int32_t* counts, int B, int N, int N_src,
                         cudaStream_t s) {
  if (B == 0) return FRODO_OK;
  FRODO_CUDA_TRY(cudaMemsetAsync(counts, 0, (size_t)B * 2 * sizeof(int32_t), s));
  if (N == 0) return FRODO_OK;
  int bx = (N + kQualThreads * 4 - 1) / (kQualThreads * 4);
  const int cap = (num_sms() * 8 + B - 1) / B;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  match_quality_kernel<<<dim3(bx, B), kQualThreads, 0, s>>>(Q_src, Q_dst, C_src, C_dst, idx,
                                                            valid_match, Q_conf, C_conf, Q_out,
                                                            valid_opt, valid_q, counts, N, N_src);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
