// Dense fp64 Cholesky solve of the assembled normal equations, pose retraction
// and the convergence test, in one kernel (no host round trip):
//   SparseBlock::solve (gn_kernels.cu:163-199; failure -> dx = 0),
//   dx = -x -> fp32 (:977, :191-192), pose_retr_kernel (:522-560),
//   ||dx||_2 < delta_thresh (:986-989).
// Hg is the (n+1) x n row-major matrix [H ; g^T]: factorising it as if g were
// one more matrix row yields y = L^-1 g in the last row for free.
#include "gn_internal.cuh"

namespace frodo {

namespace {

constexpr int NB = 32;        // block column width
constexpr int kSolveThreads = 512;
constexpr int RC = 128;       // row chunk of the trailing update
constexpr int PLD = NB + 1;   // padded leading dimension of smem panels

struct SolveSmem {
  double Ld[NB][PLD];    // factorised diagonal block
  double Pi[RC][PLD];    // panel rows of the i-chunk
  double Pj[RC][PLD];    // panel rows of the j-chunk
  double red[kSolveThreads / 32][NB];
  double xblk[NB];
  double inv_diag[NB];
  int fail;
};

// --- Sim3 retraction, gn_kernels.cu:341-455 (fp32, as the reference) ---------
__device__ void quat_comp(const float* qi, const float* qj, float* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}
__device__ void act_so3(const float* q, const float* X, float* Y) {
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  const float y0 = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  const float y1 = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  const float y2 = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
  Y[0] = y0; Y[1] = y1; Y[2] = y2;
}
__device__ void cross_inplace(const float* a, float* b) {
  const float x0 = a[1] * b[2] - a[2] * b[1];
  const float x1 = a[2] * b[0] - a[0] * b[2];
  const float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0; b[1] = x1; b[2] = x2;
}
__device__ void exp_sim3(const float* xi, float* t, float* q, float* s) {
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
__device__ void retract_pose(float* P, const float* xi) {
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

// Single CTA, right-looking blocked Cholesky on the (n+1) x n matrix in global
// memory (L2 resident), lower triangle only.
__global__ void __launch_bounds__(kSolveThreads)
solve_retract_kernel(double* __restrict__ A, float* __restrict__ Twc, float* __restrict__ dx,
                     int n, int K, int num_fix, float delta_thresh, int32_t* __restrict__ done,
                     int32_t* __restrict__ info) {
  if (done != nullptr && *done != 0) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SolveSmem& sm = *reinterpret_cast<SolveSmem*>(smem_raw);
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int rows = n + 1;  // including the rhs row
  if (tid == 0) sm.fail = 0;
  __syncthreads();

  for (int kb = 0; kb < n; kb += NB) {
    const int nb = min(NB, n - kb);
    // 1. diagonal block -> smem (lower part), factorise with one warp (lane = row)
    for (int k = tid; k < NB * NB; k += kSolveThreads) {
      const int r = k / NB, c = k % NB;
      sm.Ld[r][c] = (r < nb && c <= r) ? A[(size_t)(kb + r) * n + kb + c] : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    if (warp == 0) {
      for (int j = 0; j < nb; ++j) {
        const double d = sm.Ld[j][j];
        if (!(d > 0.0)) {
          if (lane == 0) sm.fail = 1;
          break;
        }
        const double inv = rsqrt(d);
        const double ljj = d * inv;
        __syncwarp();
        if (lane == j) {
          sm.Ld[j][j] = ljj;
          sm.inv_diag[j] = inv;
        }
        double l = 0.0;
        if (lane > j && lane < nb) {
          l = sm.Ld[lane][j] * inv;
          sm.Ld[lane][j] = l;
        }
        __syncwarp();
        if (lane > j && lane < nb) {
          for (int c = j + 1; c <= lane; ++c) sm.Ld[lane][c] -= l * sm.Ld[c][j];
        }
        __syncwarp();
      }
    }
    __syncthreads();
    if (sm.fail) break;
    // write the factorised diagonal block back (needed by the back substitution)
    for (int k = tid; k < nb * nb; k += kSolveThreads) {
      const int r = k / nb, c = k % nb;
      if (c <= r) A[(size_t)(kb + r) * n + kb + c] = sm.Ld[r][c];
    }
    // 2. panel: rows below the block (incl. rhs row): L21 = A21 L11^-T, one thread per row
    const int r0 = kb + nb;
    for (int r = r0 + tid; r < rows; r += kSolveThreads) {
      double a[NB];
      double* __restrict__ row = A + (size_t)r * n + kb;
#pragma unroll
      for (int c = 0; c < NB; ++c) a[c] = c < nb ? row[c] : 0.0;
#pragma unroll
      for (int c = 0; c < NB; ++c) {
        if (c < nb) {
          const double l = a[c] * sm.inv_diag[c];
          a[c] = l;
#pragma unroll
          for (int c2 = c + 1; c2 < NB; ++c2)
            if (c2 < nb) a[c2] -= l * sm.Ld[c2][c];
        }
      }
#pragma unroll
      for (int c = 0; c < NB; ++c)
        if (c < nb) row[c] = a[c];
    }
    __syncthreads();
    // 3. trailing update A22 -= L21 L21^T over RC x RC chunk pairs (lower part)
    const int ty = tid >> 5, tx = tid & 31;  // 16 x 32 thread grid
    for (int ci = r0; ci < rows; ci += RC) {
      const int ni = min(RC, rows - ci);
      for (int k = tid; k < RC * NB; k += kSolveThreads) {
        const int r = k / NB, c = k % NB;
        sm.Pi[r][c] = (r < ni && c < nb) ? A[(size_t)(ci + r) * n + kb + c] : 0.0;
      }
      for (int cj = r0; cj <= ci && cj < n; cj += RC) {
        const int nj = min(RC, n - cj);
        __syncthreads();
        for (int k = tid; k < RC * NB; k += kSolveThreads) {
          const int r = k / NB, c = k % NB;
          sm.Pj[r][c] = (r < nj && c < nb) ? A[(size_t)(cj + r) * n + kb + c] : 0.0;
        }
        __syncthreads();
        double acc[8][4];
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
#pragma unroll 4
        for (int k = 0; k < NB; ++k) {
          double pi[8], pj[4];
#pragma unroll
          for (int a = 0; a < 8; ++a) pi[a] = sm.Pi[ty + 16 * a][k];
#pragma unroll
          for (int b = 0; b < 4; ++b) pj[b] = sm.Pj[tx + 32 * b][k];
#pragma unroll
          for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fma(pi[a], pj[b], acc[a][b]);
        }
#pragma unroll
        for (int a = 0; a < 8; ++a) {
          const int gi = ci + ty + 16 * a;
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const int gj = cj + tx + 32 * b;
            if (gi < rows && gj < n && gj <= gi) A[(size_t)gi * n + gj] -= acc[a][b];
          }
        }
      }
      __syncthreads();
    }
    __syncthreads();
  }

  const int nopt = K - num_fix;
  if (sm.fail) {
    // not positive definite: dx = 0, poses untouched (gn_kernels.cu:193-197)
    for (int k = tid; k < nopt * 7; k += kSolveThreads) dx[k] = 0.0f;
    if (tid == 0) {
      if (info) { info[0] += 1; info[1] += 1; }
      if (done && 0.0f < delta_thresh) *done = 1;
    }
    return;
  }

  // back substitution L^T x = y, y = row n of A; x overwrites y
  double* __restrict__ y = A + (size_t)n * n;
  const int nblocks = (n + NB - 1) / NB;
  for (int bk = nblocks - 1; bk >= 0; --bk) {
    const int kb = bk * NB;
    const int nb = min(NB, n - kb);
    // partial sums over already-solved rows r >= kb+nb:  s_c = sum_r L[r][kb+c] x[r]
    double s = 0.0;
    if (lane < nb)
      for (int r = kb + nb + warp; r < n; r += kSolveThreads / 32)
        s += A[(size_t)r * n + kb + lane] * y[r];
    sm.red[warp][lane] = s;
    for (int k = tid; k < NB * NB; k += kSolveThreads) {
      const int r = k / NB, c = k % NB;
      sm.Ld[r][c] = (r < nb && c <= r) ? A[(size_t)(kb + r) * n + kb + c] : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    if (warp == 0) {
      double rhs = 0.0;
      if (lane < nb) {
        rhs = y[kb + lane];
        for (int w = 0; w < kSolveThreads / 32; ++w) rhs -= sm.red[w][lane];
      }
      for (int c = nb - 1; c >= 0; --c) {
        double xc = 0.0;
        if (lane == c) {
          xc = rhs / sm.Ld[c][c];
          sm.xblk[c] = xc;
        }
        xc = __shfl_sync(0xffffffffu, xc, c);
        if (lane < c) rhs -= sm.Ld[c][lane] * xc;
      }
      __syncwarp();
      if (lane < nb) y[kb + lane] = sm.xblk[lane];
    }
    __syncthreads();
  }

  // dx = -x (fp32), ||dx||, retraction
  __shared__ float nrm_part[kSolveThreads / 32];
  float local = 0.0f;
  for (int k = tid; k < n; k += kSolveThreads) {
    const float d = (float)(-y[k]);
    dx[k] = d;
    local += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if (lane == 0) nrm_part[warp] = local;
  __syncthreads();
  for (int k = tid; k < nopt; k += kSolveThreads) {
    float xi[7];
#pragma unroll
    for (int c = 0; c < 7; ++c) xi[c] = dx[k * 7 + c];
    float P[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) P[c] = Twc[(size_t)(k + num_fix) * 8 + c];
    retract_pose(P, xi);
#pragma unroll
    for (int c = 0; c < 8; ++c) Twc[(size_t)(k + num_fix) * 8 + c] = P[c];
  }
  if (tid == 0) {
    float tot = 0.0f;
    for (int w = 0; w < kSolveThreads / 32; ++w) tot += nrm_part[w];
    if (info) info[0] += 1;
    if (done && sqrtf(tot) < delta_thresh) *done = 1;
  }
}

}  // namespace

int launch_gn_solve_retract(double* Hg, float* Twc, float* dx, int K, int num_fix,
                            float delta_thresh, int32_t* done, int32_t* info, cudaStream_t s) {
  const int n = 7 * (K - num_fix);
  if (n <= 0) return FRODO_OK;
  static bool attr_set = false;
  if (!attr_set) {
    FRODO_CUDA_TRY(cudaFuncSetAttribute(solve_retract_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(SolveSmem)));
    attr_set = true;
  }
  solve_retract_kernel<<<1, kSolveThreads, sizeof(SolveSmem), s>>>(Hg, Twc, dx, n, K, num_fix,
                                                                  delta_thresh, done, info);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
