// EXPERIMENT (round 2, not in the product): iter_proj with TWO staged windows -- window A around
// the initial positions for lm_begin and the first gradient fetch, window B around the trial
// positions of iteration 0 (lm_iteration split into lm_propose / lm_evaluate so that the box of
// the trial positions is known before they are fetched).  Every fetch then comes from shared
// memory.  Bit-identical (all GPU tests passed), but SLOWER: 10.7 vs 11.2 G px/s on 256 pairs at
// 512x384 -- the second staging pass (three block barriers + a 40-100 KB copy per CTA) costs more
// than the ~50 scalar loads per pixel it moves from L1 to shared memory.
// iter_proj: per-pixel Levenberg-Marquardt ray projection.
// Restates matching_kernels.cu:128-284 of the reference with the exact
// rounding sequence of its nvcc 12.9 -O3 build (FMA contractions read from the
// SASS, see DESIGN.md "iter_proj arithmetic") written as explicit _rn
// intrinsics, so results are bit-identical and independent of compiler
// contraction choices.  The reference's fp64 sub-expressions are replaced by
// fp32 operations that are provably identical:
//   (1.0-du)*dv      : du is a multiple of 2^-23 in [0,1) (u >= 1), so 1-du is
//                      exact in fp32 and the exact product rounds once;
//   1.0/x -> float   : fp64 division then rounding to fp32 equals the correctly
//                      rounded fp32 reciprocal (53 >= 2*24+2, Figueroa 1995);
//   lambda *= 10.0   : exact product, single rounding.
// Only  lambda *= 0.1  (0.1 is not an fp32 value) stays in fp64.
#include <limits.h>

#include "common.cuh"
#include "ptx_async.cuh"

namespace frodo {

namespace {

__device__ __forceinline__ float sum3(float a0, float a1, float a2, float b0, float b1, float b2) {
  // a0*b0 + a1*b1 + a2*b2  ->  fma(a2,b2, fma(a0,b0, a1*b1))
  return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}

__device__ __forceinline__ float blend(float w11, float a, float w12, float b, float w21, float c,
                                       float w22, float d) {
  // w11*a + w12*b + w21*c + w22*d -> fma(w22,d, fma(w21,c, fma(w11,a, w12*b)))
  return __fmaf_rn(w22, d, __fmaf_rn(w21, c, __fmaf_rn(w11, a, __fmul_rn(w12, b))));
}

__device__ __forceinline__ float clampf(float x, float lo, float hi) {
  return fminf(fmaxf(x, lo), hi);
}

struct Weights {
  float w11, w12, w21, w22;
  int u11, v11;
};

__device__ __forceinline__ Weights bilinear_weights(float u, float v) {
  Weights w;
  w.u11 = __float2int_rd(u);
  w.v11 = __float2int_rd(v);
  const float du = __fsub_rn(u, (float)w.u11);
  const float dv = __fsub_rn(v, (float)w.v11);
  const float cu = __fsub_rn(1.0f, du);  // exact
  const float cv = __fsub_rn(1.0f, dv);  // exact
  w.w11 = __fmul_rn(du, dv);
  w.w12 = __fmul_rn(cu, dv);
  w.w21 = __fmul_rn(du, cv);
  w.w22 = __fmul_rn(cu, cv);
  return w;
}

// Levenberg-Marquardt state of one query pixel.  Consecutive iterations of the reference
// (matching_kernels.cu:162-279) recompute identical quantities: after an ACCEPTED step the next
// iteration blends the ray at exactly the trial position (same weights, same texels), so its
// ray, norm, error and cost are bit-for-bit the trial's; after a REJECTED step the position does
// not move and only lambda changes.  The state therefore carries the evaluation at (u, v) --
// error, cost, the blended d(ray_x)/dx that shares its texel quad-word with the ray -- and the
// five gradient sums, which are refreshed only after the position moved.  Same operations on the
// same inputs as the reference, so the results stay bit-identical; per iteration one 4-channel
// and (after a move) one 5-channel bilinear fetch instead of 9 + 3 channels, one sqrt/rcp pair
// instead of two.
struct Lm {
  float u, v, lambda;
  bool conv;
  Weights w;              // bilinear footprint at (u, v)
  float e0, e1, e2, cost; // ray error at (u, v)
  float gx0;              // blended channel 3 at (u, v)
  float A00, A01, A11, egx, egy;  // gradient sums at (u, v), without lambda
  bool have_grad;
};

// fetch(w, c0, c1, out): out[j - c0] = blended channel j, c0 <= j < c1, at the footprint w.
template <typename Fetch>
__device__ __forceinline__ void lm_begin(Lm& s, float tx, float ty, float tz, Fetch&& fetch) {
  s.w = bilinear_weights(s.u, s.v);
  float c[4];
  fetch(s.w, 0, 4, c);
  const float nrm = __fsqrt_rn(sum3(c[0], c[1], c[2], c[0], c[1], c[2]));
  const float inv = __frcp_rn(nrm);
  s.e0 = __fmaf_rn(c[0], inv, -tx);
  s.e1 = __fmaf_rn(c[1], inv, -ty);
  s.e2 = __fmaf_rn(c[2], inv, -tz);
  s.cost = sum3(s.e0, s.e1, s.e2, s.e0, s.e1, s.e2);
  s.gx0 = c[3];
  s.have_grad = false;
}

// First half of an iteration: the gradient sums at (u, v) if the position moved, then the
// damped step; returns the trial position.
template <typename Fetch>
__device__ __forceinline__ float2 lm_propose(Lm& s, float umax, float vmax, Fetch&& fetch) {
  if (!s.have_grad) {
    float g[5];  // gx1, gx2, gy0, gy1, gy2
    fetch(s.w, 4, 9, g);
    s.A00 = sum3(s.gx0, g[0], g[1], s.gx0, g[0], g[1]);
    s.A01 = sum3(s.gx0, g[0], g[1], g[2], g[3], g[4]);
    s.A11 = sum3(g[2], g[3], g[4], g[2], g[3], g[4]);
    s.egx = sum3(s.e0, s.e1, s.e2, s.gx0, g[0], g[1]);
    s.egy = sum3(s.e0, s.e1, s.e2, g[2], g[3], g[4]);
    s.have_grad = true;
  }
  const float A00 = __fadd_rn(s.A00, s.lambda);
  const float A11 = __fadd_rn(s.A11, s.lambda);
  const float det = __fmaf_rn(A00, A11, -__fmul_rn(s.A01, s.A01));
  const float det_inv = __frcp_rn(det);
  const float num_u = __fmaf_rn(s.A01, s.egy, -__fmul_rn(s.egx, A11));
  const float num_v = __fmaf_rn(s.A01, s.egx, -__fmul_rn(s.egy, A00));
  return make_float2(clampf(__fmaf_rn(num_u, det_inv, s.u), 1.0f, umax),
                     clampf(__fmaf_rn(num_v, det_inv, s.v), 1.0f, vmax));
}

// Second half: the ray error at the trial position, accept or reject.
template <typename Fetch>
__device__ __forceinline__ void lm_evaluate(Lm& s, float2 trial, float tx, float ty, float tz,
                                            float cost_thresh, Fetch&& fetch) {
  const float u_new = trial.x, v_new = trial.y;
  const Weights q = bilinear_weights(u_new, v_new);
  float d[4];
  fetch(q, 0, 4, d);
  const float nrm2 = __fsqrt_rn(sum3(d[0], d[1], d[2], d[0], d[1], d[2]));
  const float inv2 = __frcp_rn(nrm2);
  const float f0 = __fmaf_rn(d[0], inv2, -tx);
  const float f1 = __fmaf_rn(d[1], inv2, -ty);
  const float f2 = __fmaf_rn(d[2], inv2, -tz);
  const float new_cost = sum3(f0, f1, f2, f0, f1, f2);
  if (new_cost < s.cost) {
    s.u = u_new;
    s.v = v_new;
    s.w = q;
    s.e0 = f0; s.e1 = f1; s.e2 = f2;
    s.cost = new_cost;
    s.gx0 = d[3];
    s.have_grad = false;
    s.lambda = __double2float_rn(__dmul_rn((double)s.lambda, 0.1));
    s.conv = new_cost < cost_thresh;
  } else {
    s.lambda = __fmul_rn(s.lambda, 10.0f);
    s.conv = s.cost < cost_thresh;
  }
}

template <typename Fetch>
__device__ __forceinline__ void lm_iteration(Lm& s, float tx, float ty, float tz, float umax,
                                             float vmax, float cost_thresh, Fetch&& fetch) {
  const float2 trial = lm_propose(s, umax, vmax, fetch);
  lm_evaluate(s, trial, tx, ty, tz, cost_thresh, fetch);
}

// Result of one query pixel: the op's own outputs, or (frodo_match) the occlusion test of
// matching.py:66-75 straight from the registers -- p1 = trunc(p), valid = converged &
// (||X11[p1] - X21|| < dist_thresh), arithmetic as match_occlusion_kernel (match_glue.cu).
__device__ __forceinline__ void store_result(const Lm& s, size_t t, size_t b, int H, int W,
                                             float* __restrict__ p_new,
                                             uint8_t* __restrict__ converged,
                                             const IterProjTail& tail) {
  if (tail.X11 == nullptr) {
    reinterpret_cast<float2*>(p_new)[t] = make_float2(s.u, s.v);
    converged[t] = s.conv ? 1 : 0;
    return;
  }
  const long long px = (long long)s.u, py = (long long)s.v;  // .long(): toward zero
  reinterpret_cast<longlong2*>(tail.p1)[t] = make_longlong2(px, py);
  const long long gx = min(max(px, 0LL), (long long)W - 1), gy = min(max(py, 0LL), (long long)H - 1);
  const float* a = tail.X11 + (b * (size_t)H * W + (size_t)gy * W + (size_t)gx) * 3;
  const float* q = tail.X21 + t * 3;
  const float d0 = __fsub_rn(__ldg(a), __ldg(q));
  const float d1 = __fsub_rn(__ldg(a + 1), __ldg(q + 1));
  const float d2 = __fsub_rn(__ldg(a + 2), __ldg(q + 2));
  const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2));
  tail.valid[t] = s.conv && (__fsqrt_rn(n2) < tail.dist_thresh);
}

// One thread per query pixel; texels fetched through L1 (read-only path).
__global__ void __launch_bounds__(128)
iter_proj_kernel(const float* __restrict__ rays, const float* __restrict__ pts,
                 const float* __restrict__ p_init, float* __restrict__ p_new,
                 uint8_t* __restrict__ converged, int H, int W, int N, int max_iter,
                 float lambda_init, float cost_thresh, const IterProjTail tail) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const size_t b = blockIdx.y;
  const float* __restrict__ img = rays + b * (size_t)H * W * 9;
  const size_t t = b * (size_t)N + n;

  const float tx = __ldg(pts + t * 3 + 0);
  const float ty = __ldg(pts + t * 3 + 1);
  const float tz = __ldg(pts + t * 3 + 2);
  const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + t);
  const float umax = (float)(W - 2), vmax = (float)(H - 2);
  Lm s;
  s.u = clampf(p0.x, 1.0f, umax);
  s.v = clampf(p0.y, 1.0f, vmax);
  s.lambda = lambda_init;
  s.conv = false;
  auto fetch_global = [&](const Weights& w, int c0, int c1, float* out) {
    const float* r22 = img + ((size_t)w.v11 * W + w.u11) * 9;  // (v11, u11); +9: u11+1; +9W: v11+1
    const float* r12 = r22 + (size_t)W * 9;
#pragma unroll
    for (int j = 0; j < 9; ++j)
      if (j >= c0 && j < c1)
        out[j - c0] = blend(w.w11, __ldg(r12 + 9 + j), w.w12, __ldg(r12 + j), w.w21,
                            __ldg(r22 + 9 + j), w.w22, __ldg(r22 + j));
  };
  if (max_iter > 0) lm_begin(s, tx, ty, tz, fetch_global);
  for (int it = 0; it < max_iter; ++it)
    lm_iteration(s, tx, ty, tz, umax, vmax, cost_thresh, fetch_global);
  store_result(s, t, b, H, W, p_new, converged, tail);  // the tail needs N == H * W
}

// ---------------------------------------------------------------------------------------
// Window-staged variant.  ncu on the kernel above: 31 sectors and 3.3 L1 wavefronts per
// scalar texel load (the 36-byte AoS texel stride puts every lane in its own sector), 77.8 M
// L1 wavefronts per 8 pairs = 93 % of its run time, top stall lg_throttle.  Shared memory
// serves the same 4-byte loads at one wavefront per warp (lane stride 9 words is coprime
// with the 32 banks).  A CTA owns a 32x16 patch of query pixels.  LM iteration 0 runs from
// global memory (it moves the points from the initial guess to the neighbourhood of the
// solution, mean 16 px on the synthetic pairs); then the bounding box of the patch's current
// positions plus a margin is copied into shared memory with one cp.async.bulk per box row and
// iterations 1.. fetch from there; a 2x2 footprint outside the box falls back to global
// memory, so results never depend on the staging.  Arithmetic is shared with the kernel above.
constexpr int kWinPatchW = 32, kWinPatchH = 16, kWinThreads = kWinPatchW * kWinPatchH;
constexpr int kWinMargin = 4;
constexpr int kWinSmemBytes = 108 * 1024;  // 2 CTAs per SM

__global__ void __launch_bounds__(kWinThreads, 2)
iter_proj_window_kernel(const float* __restrict__ rays, const float* __restrict__ pts,
                        const float* __restrict__ p_init, float* __restrict__ p_new,
                        uint8_t* __restrict__ converged, int H, int W, int max_iter,
                        float lambda_init, float cost_thresh, const IterProjTail tail) {
  extern __shared__ __align__(128) float win[];
  __shared__ uint64_t bar;
  __shared__ int box[4];
  const int tid = threadIdx.x;
  const int x = blockIdx.x * kWinPatchW + (tid & (kWinPatchW - 1));
  const int y = blockIdx.y * kWinPatchH + (tid / kWinPatchW);
  const size_t b = blockIdx.z;
  const bool active = x < W && y < H;
  const float* __restrict__ img = rays + b * (size_t)H * W * 9;
  const size_t t = b * (size_t)H * W + (size_t)y * W + x;
  if (tid == 0) {
    box[0] = INT_MAX; box[1] = INT_MIN; box[2] = INT_MAX; box[3] = INT_MIN;
    ptx::mbar_init(&bar, 1);
    ptx::mbar_fence_init();
  }
  float tx = 0.f, ty = 0.f, tz = 0.f;
  const float umax = (float)(W - 2), vmax = (float)(H - 2);
  Lm s;
  s.u = s.v = 1.0f;
  s.lambda = lambda_init;
  s.conv = false;
  if (active) {
    tx = __ldg(pts + t * 3 + 0);
    ty = __ldg(pts + t * 3 + 1);
    tz = __ldg(pts + t * 3 + 2);
    const float2 p0 = __ldg(reinterpret_cast<const float2*>(p_init) + t);
    s.u = clampf(p0.x, 1.0f, umax);
    s.v = clampf(p0.y, 1.0f, vmax);
  }
  auto fetch_global = [&](const Weights& w, int c0, int c1, float* out) {
    const float* r22 = img + ((size_t)w.v11 * W + w.u11) * 9;
    const float* r12 = r22 + (size_t)W * 9;
#pragma unroll
    for (int j = 0; j < 9; ++j)
      if (j >= c0 && j < c1)
        out[j - c0] = blend(w.w11, __ldg(r12 + 9 + j), w.w12, __ldg(r12 + j), w.w21,
                            __ldg(r22 + 9 + j), w.w22, __ldg(r22 + j));
  };
  // Window geometry (texel box [tu0, tu0+TW) x [tv0, tv0+TH) in shared memory), set by stage().
  int tu0 = 0, tv0 = 0, TW = 0, TH = 0;
  uint32_t bar_parity = 0;
  // Stages the bounding box of the footprints whose top-left texels are (pu, pv) over the
  // active threads of the CTA, plus a margin.  Every thread of the CTA must call it; on return
  // the window is in shared memory (or empty: TW = 0) and all earlier reads of it are over.
  auto stage = [&](int pu, int pv) {
    __syncthreads();  // earlier window reads are over; box[] of the previous call consumed
    if (tid == 0) {
      box[0] = INT_MAX; box[1] = INT_MIN; box[2] = INT_MAX; box[3] = INT_MIN;
    }
    __syncthreads();
    int mnu = INT_MAX, mxu = INT_MIN, mnv = INT_MAX, mxv = INT_MIN;
    if (active) {
      mnu = mxu = pu;
      mnv = mxv = pv;
    }
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
    __syncthreads();
    tu0 = tv0 = TW = TH = 0;
    if (box[0] > box[1]) return;
    constexpr int cap = kWinSmemBytes / 36;  // texels
    tu0 = max(box[0] - kWinMargin, 0) & ~3;                  // 16-byte aligned row starts
    const int tu1 = min(box[1] + 1 + kWinMargin, W - 1);     // +1: right texel of the footprint
    tv0 = max(box[2] - kWinMargin, 0);
    const int tv1 = min(box[3] + 1 + kWinMargin, H - 1);
    TW = (tu1 - tu0 + 1 + 3) & ~3;
    if (tu0 + TW > W) TW = W - tu0;  // W % 4 == 0 keeps this a multiple of 4
    TH = tv1 - tv0 + 1;
    if (TW > (cap & ~3)) TW = cap & ~3;
    if (TW * TH > cap) {
      const int keep = cap / TW;
      tv0 += (TH - keep) / 2;
      TH = keep;
    }
    if (tid < 32) {
      const uint32_t row_bytes = (uint32_t)TW * 36u;
      if (tid == 0) ptx::mbar_arrive_expect_tx(&bar, row_bytes * (uint32_t)TH);
      __syncwarp();
      for (int r = tid; r < TH; r += 32)
        ptx::bulk_g2s(win + (size_t)r * TW * 9, img + ((size_t)(tv0 + r) * W + tu0) * 9, row_bytes,
                      &bar);
    }
    ptx::mbar_wait(&bar, bar_parity);
    bar_parity ^= 1u;
  };
  auto fetch_window = [&](const Weights& w, int c0, int c1, float* out) {
    const int ru = w.u11 - tu0, rv = w.v11 - tv0;
    if (ru >= 0 && ru + 1 < TW && rv >= 0 && rv + 1 < TH) {
      const float* r22 = win + (rv * TW + ru) * 9;
      const float* r12 = r22 + TW * 9;
#pragma unroll
      for (int j = 0; j < 9; ++j)
        if (j >= c0 && j < c1)
          out[j - c0] = blend(w.w11, r12[9 + j], w.w12, r12[j], w.w21, r22[9 + j], w.w22, r22[j]);
    } else {
      fetch_global(w, c0, c1, out);
    }
  };
  if (max_iter <= 0) {
    if (active) store_result(s, t, b, H, W, p_new, converged, tail);
    return;
  }
  // Window A: around the initial positions -- the first evaluation and the first gradient.
  stage(__float2int_rd(s.u), __float2int_rd(s.v));
  float2 trial = make_float2(s.u, s.v);
  if (active) {
    lm_begin(s, tx, ty, tz, fetch_window);
    trial = lm_propose(s, umax, vmax, fetch_window);
  }
  // Window B: around the trial positions of iteration 0 (the step that carries the points from
  // the initial guess to the neighbourhood of the solution); iterations 1.. stay inside it.
  stage(__float2int_rd(trial.x), __float2int_rd(trial.y));
  if (active) {
    lm_evaluate(s, trial, tx, ty, tz, cost_thresh, fetch_window);
    for (int it = 1; it < max_iter; ++it)
      lm_iteration(s, tx, ty, tz, umax, vmax, cost_thresh, fetch_window);
    store_result(s, t, b, H, W, p_new, converged, tail);
  }
}

}  // namespace

int launch_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_new,
                     uint8_t* converged, int B, int H, int W, int N, int max_iter,
                     float lambda_init, float cost_thresh, cudaStream_t s,
                     const IterProjTail* tail_in) {
  if (B == 0 || N == 0) return FRODO_OK;
  IterProjTail tail = {};
  if (tail_in != nullptr) {
    tail = *tail_in;
    if (N != H * W) {
      set_error("iter_proj: the occlusion tail needs N == H * W");
      return FRODO_ERR_ARG;
    }
  }
  const bool aligned = (reinterpret_cast<uintptr_t>(rays) & 15) == 0 &&
                       (reinterpret_cast<uintptr_t>(p_init) & 7) == 0 &&
                       (reinterpret_cast<uintptr_t>(p_new) & 7) == 0;
  if (N == H * W && (W % 4) == 0 && W >= 4 && H >= 3 && aligned && B <= 65535) {
    static PerDeviceOnce once;
    int dev;
    if (once.needed(&dev)) {
      FRODO_CUDA_TRY(cudaFuncSetAttribute(iter_proj_window_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          kWinSmemBytes));
      once.mark(dev);
    }
    dim3 wgrid((W + kWinPatchW - 1) / kWinPatchW, (H + kWinPatchH - 1) / kWinPatchH, B);
    iter_proj_window_kernel<<<wgrid, kWinThreads, kWinSmemBytes, s>>>(
        rays, pts, p_init, p_new, converged, H, W, max_iter, lambda_init, cost_thresh, tail);
    FRODO_CUDA_TRY(cudaGetLastError());
    return FRODO_OK;
  }
  dim3 grid((N + 127) / 128, B);
  iter_proj_kernel<<<grid, 128, 0, s>>>(rays, pts, p_init, p_new, converged, H, W, N, max_iter,
                                        lambda_init, cost_thresh, tail);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
