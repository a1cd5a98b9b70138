// Pose update shared by the Cholesky kernels: Sim3 retraction (gn_kernels.cu:341-455,
// pose_retr_kernel :522-560), dx = -x in fp32 (:977, :191-192), the convergence test
// (:986-989) and the not-positive-definite path (:193-197).
#pragma once
#include "gn_internal.cuh"

namespace frodo {
namespace pose_update {

// --- Sim3 retraction, gn_kernels.cu:341-455 (fp32, as the reference) ---------
__device__ inline void quat_comp(const float* qi, const float* qj, float* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}
__device__ inline void act_so3(const float* q, const float* X, float* Y) {
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const float y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const float y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}
__device__ inline void cross_inplace(const float* a, float* b) {
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}
__device__ inline void exp_sim3(const float* xi, float* t, float* q, float* s) {
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float phi[3] = {xi[3], xi[4], xi[5]};
  const float sigma = xi[6];
  const float scale = expf(sigma);
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta = sqrtf(theta_sq);
  float imag, real;
  if (theta_sq < 1e-6f) {  // expSO3 :348-356 (fp64 compare and series in the reference)
    const double t2 = theta_sq, t4 = (double)(theta_sq * theta_sq);
    imag = (float)(0.5 - (1.0 / 48.0) * t2 + (1.0 / 3840.0) * t4);
    real = (float)(1.0 - (1.0 / 8.0) * t2 + (1.0 / 384.0) * t4);
  } else {
    imag = sinf(0.5f * theta) / theta;
    real = cosf(0.5f * theta);
  }
  q[0] = imag * phi[0]; q[1] = imag * phi[1]; q[2] = imag * phi[2]; q[3] = real;
  s[0] = scale;
  float A, B, C;
  const double EPSd = 1e-6;
  if ((double)fabsf(sigma) < EPSd) {
    C = 1.0f;
    if ((double)fabsf(theta) < EPSd) {
      A = 0.5f;
      B = (float)(1.0 / 6.0);
    } else {
      A = (1.0f - cosf(theta)) / theta_sq;
      B = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - 1.0f) / sigma;
    if ((double)fabsf(theta) < EPSd) {
      const float sigma_sq = sigma * sigma;
      A = ((sigma - 1.0f) * scale + 1.0f) / sigma_sq;
      B = (scale * 0.5f * sigma_sq + scale - 1.0f - sigma * scale) / (sigma_sq * sigma);
    } else {
      const float a = scale * sinf(theta);
      const float b = scale * cosf(theta);
      const float c = theta_sq + sigma * sigma;
      A = (a * sigma + (1.0f - b) * theta) / (theta * c);
      B = (C - ((b - 1.0f) * sigma + a * theta) / c) / theta_sq;
    }
  }
  t[0] = C * tau[0]; t[1] = C * tau[1]; t[2] = C * tau[2];
  cross_inplace(phi, tau);
  t[0] += A * tau[0]; t[1] += A * tau[1]; t[2] += A * tau[2];
  cross_inplace(phi, tau);
  t[0] += B * tau[0]; t[1] += B * tau[1]; t[2] += B * tau[2];
}
__device__ inline void retract_pose(float* P, const float* xi) {
  float dt[3], dq[4], ds[1];
  exp_sim3(xi, dt, dq, ds);
  const float t[3] = {P[0], P[1], P[2]};
  const float q[4] = {P[3], P[4], P[5], P[6]};
  float q1[4], t1[3];
  quat_comp(dq, q, q1);
  act_so3(dq, t, t1);
  P[0] = t1[0] * ds[0] + dt[0];
  P[1] = t1[1] * ds[0] + dt[1];
  P[2] = t1[2] * ds[0] + dt[2];
  P[3] = q1[0]; P[4] = q1[1]; P[5] = q1[2]; P[6] = q1[3];
  P[7] = ds[0] * P[7];
}

// dx = -x (fp32, :977, :191-192), ||dx||_2 < delta_thresh (:986-989), pose
// retraction (pose_retr_kernel :522-560).  Called by all THREADS threads of one CTA.
// pose_rows (may be NULL): row of Twc that holds keyframe r of the linear system's ordering
// (keyframe arena, graph.py); NULL = identity, the reference's dense [K,8] layout.
template <int THREADS>
__device__ inline void finish_update(const double* x, float* __restrict__ Twc, float* __restrict__ dx,
                              int n, int nopt, int num_fix, float delta_thresh,
                              int32_t* __restrict__ done, int32_t* __restrict__ info,
                              const int32_t* __restrict__ pose_rows = nullptr) {
  __shared__ float nrm_part[THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float local = 0.0f;
  for (int k = tid; k < n; k += THREADS) {
    const float d = (float)(-x[k]);
    dx[k] = d;
    local += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if (lane == 0) nrm_part[warp] = local;
  __syncthreads();
  for (int k = tid; k < nopt; k += THREADS) {
    float xi[7];
#pragma unroll
    for (int c = 0; c < 7; ++c) xi[c] = dx[k * 7 + c];
    float P[8];
    const size_t row = pose_rows ? (size_t)pose_rows[k + num_fix] : (size_t)(k + num_fix);
#pragma unroll
    for (int c = 0; c < 8; ++c) P[c] = Twc[row * 8 + c];
    retract_pose(P, xi);
#pragma unroll
    for (int c = 0; c < 8; ++c) Twc[row * 8 + c] = P[c];
  }
  if (tid == 0) {
    float tot = 0.0f;
    for (int w = 0; w < THREADS / 32; ++w) tot += nrm_part[w];
    if (info) info[0] += 1;
    if (done && sqrtf(tot) < delta_thresh) *done = 1;
  }
}

// Not positive definite: dx = 0, poses untouched (gn_kernels.cu:193-197).
template <int THREADS>
__device__ inline void fail_update(float* __restrict__ dx, int nopt, float delta_thresh,
                            int32_t* __restrict__ done, int32_t* __restrict__ info) {
  for (int k = threadIdx.x; k < nopt * 7; k += THREADS) dx[k] = 0.0f;
  if (threadIdx.x == 0) {
    if (info) { info[0] += 1; info[1] += 1; }
    if (done && 0.0f < delta_thresh) *done = 1;
  }
}

}  // namespace pose_update
}  // namespace frodo
