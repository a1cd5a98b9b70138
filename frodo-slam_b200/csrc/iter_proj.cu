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
#include "common.cuh"

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

// One thread per query pixel; texels fetched through L1 (read-only path).
__global__ void __launch_bounds__(128)
iter_proj_kernel(const float* __restrict__ rays, const float* __restrict__ pts,
                 const float* __restrict__ p_init, float* __restrict__ p_new,
                 uint8_t* __restrict__ converged, int H, int W, int N, int max_iter,
                 float lambda_init, float cost_thresh) {
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
  float u = clampf(p0.x, 1.0f, umax);
  float v = clampf(p0.y, 1.0f, vmax);
  float lambda = lambda_init;
  bool conv = false;

  for (int it = 0; it < max_iter; ++it) {
    const Weights w = bilinear_weights(u, v);
    const float* r22 = img + ((size_t)w.v11 * W + w.u11) * 9;  // (v11,   u11)
    const float* r21 = r22 + 9;                                 // (v11,   u11+1)
    const float* r12 = r22 + (size_t)W * 9;                     // (v11+1, u11)
    const float* r11 = r12 + 9;                                 // (v11+1, u11+1)
    float c[9];
#pragma unroll
    for (int j = 0; j < 9; ++j)
      c[j] = blend(w.w11, __ldg(r11 + j), w.w12, __ldg(r12 + j), w.w21, __ldg(r21 + j), w.w22,
                   __ldg(r22 + j));
    const float nrm = __fsqrt_rn(sum3(c[0], c[1], c[2], c[0], c[1], c[2]));
    const float inv = __frcp_rn(nrm);
    const float e0 = __fmaf_rn(c[0], inv, -tx);
    const float e1 = __fmaf_rn(c[1], inv, -ty);
    const float e2 = __fmaf_rn(c[2], inv, -tz);
    const float cost = sum3(e0, e1, e2, e0, e1, e2);
    float A00 = sum3(c[3], c[4], c[5], c[3], c[4], c[5]);
    const float A01 = sum3(c[3], c[4], c[5], c[6], c[7], c[8]);
    float A11 = sum3(c[6], c[7], c[8], c[6], c[7], c[8]);
    const float egx = sum3(e0, e1, e2, c[3], c[4], c[5]);
    const float egy = sum3(e0, e1, e2, c[6], c[7], c[8]);
    A00 = __fadd_rn(A00, lambda);
    A11 = __fadd_rn(A11, lambda);
    const float det = __fmaf_rn(A00, A11, -__fmul_rn(A01, A01));
    const float det_inv = __frcp_rn(det);
    const float num_u = __fmaf_rn(A01, egy, -__fmul_rn(egx, A11));
    const float num_v = __fmaf_rn(A01, egx, -__fmul_rn(egy, A00));
    const float u_new = clampf(__fmaf_rn(num_u, det_inv, u), 1.0f, umax);
    const float v_new = clampf(__fmaf_rn(num_v, det_inv, v), 1.0f, vmax);

    const Weights q = bilinear_weights(u_new, v_new);
    const float* s22 = img + ((size_t)q.v11 * W + q.u11) * 9;
    const float* s21 = s22 + 9;
    const float* s12 = s22 + (size_t)W * 9;
    const float* s11 = s12 + 9;
    float d[3];
#pragma unroll
    for (int j = 0; j < 3; ++j)
      d[j] = blend(q.w11, __ldg(s11 + j), q.w12, __ldg(s12 + j), q.w21, __ldg(s21 + j), q.w22,
                   __ldg(s22 + j));
    const float nrm2 = __fsqrt_rn(sum3(d[0], d[1], d[2], d[0], d[1], d[2]));
    const float inv2 = __frcp_rn(nrm2);
    const float f0 = __fmaf_rn(d[0], inv2, -tx);
    const float f1 = __fmaf_rn(d[1], inv2, -ty);
    const float f2 = __fmaf_rn(d[2], inv2, -tz);
    const float new_cost = sum3(f0, f1, f2, f0, f1, f2);

    if (new_cost < cost) {
      u = u_new;
      v = v_new;
      lambda = __double2float_rn(__dmul_rn((double)lambda, 0.1));
      conv = new_cost < cost_thresh;
    } else {
      lambda = __fmul_rn(lambda, 10.0f);
      conv = cost < cost_thresh;
    }
  }
  reinterpret_cast<float2*>(p_new)[t] = make_float2(u, v);
  converged[t] = conv ? 1 : 0;
}

}  // namespace

int launch_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_new,
                     uint8_t* converged, int B, int H, int W, int N, int max_iter,
                     float lambda_init, float cost_thresh, cudaStream_t s) {
  if (B == 0 || N == 0) return FRODO_OK;
  dim3 grid((N + 127) / 128, B);
  iter_proj_kernel<<<grid, 128, 0, s>>>(rays, pts, p_init, p_new, converged, H, W, N, max_iter,
                                        lambda_init, cost_thresh);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
