// Dense fp64 Cholesky solve of the assembled normal equations, pose retraction
// and the convergence test, in one kernel (no host round trip):
//   SparseBlock::solve (gn_kernels.cu:163-199; failure -> dx = 0),
//   dx = -x -> fp32 (:977, :191-192), pose_retr_kernel (:522-560),
//   ||dx||_2 < delta_thresh (:986-989).
// Hg is the (n+1) x n row-major matrix [H ; g^T]: factorising it as if g were
// one more matrix row yields y = L^-1 g in the last row for free.
#include <cooperative_groups.h>

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

// dx = -x (fp32, :977, :191-192), ||dx||_2 < delta_thresh (:986-989), pose
// retraction (pose_retr_kernel :522-560).  Called by all THREADS threads of one CTA.
template <int THREADS>
__device__ void finish_update(const double* x, float* __restrict__ Twc, float* __restrict__ dx,
                              int n, int nopt, int num_fix, float delta_thresh,
                              int32_t* __restrict__ done, int32_t* __restrict__ info) {
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
#pragma unroll
    for (int c = 0; c < 8; ++c) P[c] = Twc[(size_t)(k + num_fix) * 8 + c];
    retract_pose(P, xi);
#pragma unroll
    for (int c = 0; c < 8; ++c) Twc[(size_t)(k + num_fix) * 8 + c] = P[c];
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
__device__ void fail_update(float* __restrict__ dx, int nopt, float delta_thresh,
                            int32_t* __restrict__ done, int32_t* __restrict__ info) {
  for (int k = threadIdx.x; k < nopt * 7; k += THREADS) dx[k] = 0.0f;
  if (threadIdx.x == 0) {
    if (info) { info[0] += 1; info[1] += 1; }
    if (done && 0.0f < delta_thresh) *done = 1;
  }
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
    fail_update<kSolveThreads>(dx, nopt, delta_thresh, done, info);
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

  finish_update<kSolveThreads>(y, Twc, dx, n, nopt, num_fix, delta_thresh, done, info);
}


// ============================================================================
// Cluster-resident tiled Cholesky (the fast path).
//
// The lower triangle of the (n+1) x n system [H ; g^T] is cut into 32x32 fp64
// tiles that live in the shared memory of a thread-block cluster (tile t =
// i(i+1)/2 + j is owned by CTA t % C, slot t / C), so the whole factorisation
// runs without touching global memory.  Per tile column k:
//   a. the owner of (k,k) factors it (warp 0, rows in registers, left-looking)
//      while warp 1 builds W = L_kk^-1 in lock step; W is pushed to every CTA;
//   b. cluster barrier;
//   c. owners of (i,k), i > k, form L_ik = A_ik W^T (a 32^3 tile product) and
//      push it into the panel buffer of every CTA over distributed shared memory;
//   d. cluster barrier;
//   e. every CTA applies A_ij -= L_ik L_jk^T to the tiles it owns (j > k).
// The rhs row g^T rides along as row n, so y = L^-1 g falls out of the panel
// solves; the final back substitution (CTA 0, tiles read back from L2) uses the
// stored W_k, so it is matrix-vector work only.
// Tiles are column-major with a 16-byte XOR swizzle (swz) so that the 4x4
// register-blocked tile product and its read-modify-write are bank-conflict free.
// ============================================================================
namespace cg = cooperative_groups;

constexpr int TB = 32;            // tile edge
constexpr int kTile = TB * TB;    // doubles per tile (8 KB)
constexpr int kCT = 256;          // threads per CTA (the fp64 pipe saturates with 8 warps)
constexpr int kGroups = kCT / 64; // 64-thread groups, one tile product each
constexpr int kWarps = kCT / 32;
constexpr int kZ = 1024 / kCT;    // z registers per thread in the back substitution (Tc <= 32)
constexpr int kMaxOwn = 32;       // tile slots per CTA (bounded by shared memory anyway)
constexpr int kLsLd = 33;         // leading dimension of the row-major scratch of the diagonal factor

__device__ __forceinline__ int swz(int r, int c) {
  return (c << 5) + ((((r >> 1) ^ ((c >> 1) & 7)) << 1) | (r & 1));
}

__device__ __forceinline__ void tile_decode(int t, int& i, int& j) {
  int a = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= t) ++a;
  while (a * (a + 1) / 2 > t) --a;
  i = a;
  j = t - a * (a + 1) / 2;
}

__device__ __forceinline__ void group_barrier(int g) {
  asm volatile("bar.sync %0, 64;" ::"r"(g + 1) : "memory");
}

// acc[a][b] += sum_kk A(R(a), kk) * B(C(b), kk) for two swizzled column-major tiles;
// thread (ty,tx) of an 8x8 group owns rows R(a) = 2ty + (a&1) + 16(a>>1) and columns
// C(b) = 2tx + (b&1) + 16(b>>1).
__device__ __forceinline__ void tile_mma(const double* __restrict__ A, const double* __restrict__ B,
                                         int ty, int tx, double (&acc)[4][4]) {
#pragma unroll 8
  for (int kk = 0; kk < TB; ++kk) {
    const int s = (kk >> 1) & 7;
    const double2 a0 = *reinterpret_cast<const double2*>(A + (kk << 5) + ((ty ^ s) << 1));
    const double2 a1 = *reinterpret_cast<const double2*>(A + (kk << 5) + (((ty + 8) ^ s) << 1));
    const double2 b0 = *reinterpret_cast<const double2*>(B + (kk << 5) + ((tx ^ s) << 1));
    const double2 b1 = *reinterpret_cast<const double2*>(B + (kk << 5) + (((tx + 8) ^ s) << 1));
    const double av[4] = {a0.x, a0.y, a1.x, a1.y};
    const double bv[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
  }
}

// Offset of the element pair (R(a), R(a)+1 ; C(b)) of thread (ty,tx), a in {0,2}.
__device__ __forceinline__ int out_pair_off(int ty, int tx, int a, int b) {
  const int c = 2 * tx + (b & 1) + 16 * (b >> 1);
  return (c << 5) + (((ty + 8 * (a >> 1)) ^ tx) << 1);
}

// Warp 0 of the diagonal owner: in-register left-looking Cholesky of one tile.
// Lane r holds row r.  Writes L row-major into Ls (for warp 1), 1/L_jj into invd,
// and on the tile that also carries the rhs row (lane nv) its y values to ytail.
__device__ void diag_factor_warp(const double* __restrict__ tile, double* __restrict__ Ls,
                                 double* __restrict__ invd, volatile int* progress,
                                 volatile int* fail, int nv, int rhs_row,
                                 double* __restrict__ ytail) {
  const int r = threadIdx.x & 31;
  double a[TB];
#pragma unroll
  for (int c = 0; c < TB; ++c) a[c] = tile[swz(r, c)];
#pragma unroll
  for (int j = 0; j < TB; ++j) {
    if (j < nv) {
      double s = a[j];
#pragma unroll
      for (int m = 0; m + 1 < j; ++m) s = fma(-a[m], Ls[j * kLsLd + m], s);
      if (j >= 1) {
        const double bl = __shfl_sync(0xffffffffu, a[j - 1], j);
        s = fma(-a[j - 1], bl, s);
      }
      const double d = __shfl_sync(0xffffffffu, s, j);
      if (!(d > 0.0)) {
        if (r == 0) *fail = 1;
        break;
      }
      const double inv = rsqrt(d);
      const double l = (r == j) ? d * inv : s * inv;
      a[j] = (r >= j) ? l : 0.0;
      Ls[r * kLsLd + j] = a[j];
      if (r == j) invd[j] = inv;
      __syncwarp();
      if (r == 0) {
        __threadfence_block();
        *progress = j + 1;
      }
    }
  }
  if (rhs_row >= 0 && r == rhs_row) {
#pragma unroll
    for (int c = 0; c < TB; ++c) ytail[c] = (c < nv) ? a[c] : 0.0;
  }
}

// Warp 1 of the diagonal owner: W = L^-1 (lower triangular), lane c holds column c.
__device__ void diag_inverse_warp(double* __restrict__ Wt, const double* __restrict__ Ls,
                                  const double* __restrict__ invd, volatile int* progress,
                                  volatile int* fail, int nv) {
  const int c = threadIdx.x & 31;
  double w[TB];
#pragma unroll
  for (int j = 0; j < TB; ++j) {
    if (j < nv) {
      while (*progress < j + 1 && *fail == 0) {
      }
      if (*fail != 0) break;
      __threadfence_block();
      double t = (j == c) ? 1.0 : 0.0;
#pragma unroll
      for (int m = 0; m < j; ++m) t = fma(-Ls[j * kLsLd + m], w[m], t);
      w[j] = t * invd[j];
    } else {
      w[j] = (j == c) ? 1.0 : 0.0;
    }
  }
#pragma unroll
  for (int j = 0; j < TB; j += 2)
    *reinterpret_cast<double2*>(Wt + swz(j, c)) = make_double2(w[j], w[j + 1]);
}

struct ClusterArgs {
  const double* Hg;   // [(n+1) x n] row-major, rhs in row n
  double* gt;         // global tile scratch: T(T+1)/2 tiles + 32 doubles (ytail)
  float* Twc;
  float* dx;
  int32_t* done;
  int32_t* info;
  int n, K, num_fix, T, Tc, max_own, npanel;
  float delta_thresh;
};

__global__ void __launch_bounds__(kCT, 1) chol_cluster_kernel(ClusterArgs p) {
  if (p.done != nullptr && *p.done != 0) return;  // uniform over the cluster
  cg::cluster_group cluster = cg::this_cluster();
  const int C = (int)cluster.num_blocks();
  const int rank = (int)cluster.block_rank();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* own = reinterpret_cast<double*>(smem_raw);
  double* panel = own + (size_t)p.max_own * kTile;
  double* wbuf = panel + (size_t)p.npanel * kTile;
  __shared__ int s_ti[kMaxOwn], s_tj[kMaxOwn];
  __shared__ int s_nown, s_fail, s_progress;
  __shared__ double s_invd[TB];
  __shared__ double s_z[TB], s_x[TB];

  const int tid = threadIdx.x;
  const int n = p.n, T = p.T, Tc = p.Tc;
  const int ntiles = T * (T + 1) / 2;
  const int g = tid >> 6, tg = tid & 63, ty = tg >> 3, tx = tg & 7;

  if (tid == 0) {
    s_fail = 0;
    s_progress = 0;
    int cnt = 0;
    for (int t = rank; t < ntiles; t += C) {
      int i, j;
      tile_decode(t, i, j);
      s_ti[cnt] = i;
      s_tj[cnt] = j;
      ++cnt;
    }
    s_nown = cnt;
  }
  __syncthreads();
  const int nown = s_nown;

  // load the owned tiles (transposing: global rows are contiguous in c)
  for (int sl = 0; sl < nown; ++sl) {
    const int i = s_ti[sl], j = s_tj[sl];
    double* tile = own + (size_t)sl * kTile;
    for (int e = tid; e < kTile; e += kCT) {
      const int r = e >> 5, c = e & 31;
      const int R = TB * i + r, Cg = TB * j + c;
      double v = (R == Cg) ? 1.0 : 0.0;
      if (R <= n && Cg < n) v = __ldcg(p.Hg + (size_t)R * n + Cg);
      tile[swz(r, c)] = v;
    }
  }
  cluster.sync();

  const int rl = n - TB * (T - 1);  // local index of the rhs row inside tile row T-1
  double* ytail = p.gt + (size_t)ntiles * kTile;

  for (int k = 0; k < Tc; ++k) {
    const int tkk = k * (k + 1) / 2 + k;
    const int nv = min(TB, n - TB * k);
    // ---- a. diagonal tile: factor + inverse on its owner ------------------
    if (tkk % C == rank) {
      double* tile = own + (size_t)(tkk / C) * kTile;
      double* Ls = panel;  // the panel buffer is idle here (npanel >= 2 tiles >= 33*32 doubles)
      if (tid == 0) s_progress = 0;
      __syncthreads();
      if (tid < 32) {
        diag_factor_warp(tile, Ls, s_invd, &s_progress, &s_fail, nv,
                         (k == T - 1) ? rl : -1, ytail);
      } else if (tid < 64) {
        diag_inverse_warp(wbuf, Ls, s_invd, &s_progress, &s_fail, nv);
      }
      __syncthreads();
      if (s_fail) {
        if (tid < C && tid != rank) *cluster.map_shared_rank(&s_fail, tid) = 1;
      } else {
        // W replaces L_kk in the owner's slot and goes to every CTA's wbuf
        for (int e = 2 * tid; e < kTile; e += 2 * kCT) {
          const double2 v = *reinterpret_cast<const double2*>(wbuf + e);
          *reinterpret_cast<double2*>(tile + e) = v;
          for (int d = 0; d < C; ++d)
            if (d != rank)
              *reinterpret_cast<double2*>(cluster.map_shared_rank(wbuf, d) + e) = v;
        }
      }
    }
    cluster.sync();  // b
    if (s_fail) break;
    // ---- c. panel: L_ik = A_ik W^T on the owners, pushed to every CTA -----
    {
      int q = 0;
      for (int sl = 0; sl < nown; ++sl) {
        if (s_tj[sl] != k || s_ti[sl] == k) continue;
        if ((q++ % kGroups) != g) continue;
        double* tile = own + (size_t)sl * kTile;
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        tile_mma(tile, wbuf, ty, tx, acc);
        group_barrier(g);  // every thread of the group is done reading the tile
        const int ps = s_ti[sl] - k - 1;
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
          for (int a = 0; a < 4; a += 2) {
            const int off = out_pair_off(ty, tx, a, b);
            const double2 v = make_double2(acc[a][b], acc[a + 1][b]);
            *reinterpret_cast<double2*>(tile + off) = v;
            for (int d = 0; d < C; ++d)
              *reinterpret_cast<double2*>(cluster.map_shared_rank(panel, d) +
                                          (size_t)ps * kTile + off) = v;
          }
      }
    }
    cluster.sync();  // d
    // ---- e. trailing update of the owned tiles ----------------------------
    {
      int q = 0;
      for (int sl = 0; sl < nown; ++sl) {
        const int i = s_ti[sl], j = s_tj[sl];
        if (j <= k) continue;
        if ((q++ % kGroups) != g) continue;
        double* tile = own + (size_t)sl * kTile;
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        tile_mma(panel + (size_t)(i - k - 1) * kTile, panel + (size_t)(j - k - 1) * kTile, ty, tx,
                 acc);
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
          for (int a = 0; a < 4; a += 2) {
            double2* ptr = reinterpret_cast<double2*>(tile + out_pair_off(ty, tx, a, b));
            double2 v = *ptr;
            v.x -= acc[a][b];
            v.y -= acc[a + 1][b];
            *ptr = v;
          }
      }
    }
    __syncthreads();  // the next diagonal owner must see its updated tile
  }

  const int failed = s_fail;
  // factor tiles (L_ik, and W_k on the diagonal) -> global scratch for the back substitution
  if (!failed) {
    for (int sl = 0; sl < nown; ++sl) {
      const int t = sl * C + rank;
      const double2* src = reinterpret_cast<const double2*>(own + (size_t)sl * kTile);
      double2* dst = reinterpret_cast<double2*>(p.gt + (size_t)t * kTile);
      for (int e = tid; e < kTile / 2; e += kCT) dst[e] = src[e];
    }
    __threadfence();
  }
  cluster.sync();  // also: no CTA exits while its shared memory may still be written remotely
  if (rank != 0) return;

  const int nopt = p.K - p.num_fix;
  if (failed) {
    fail_update<kCT>(p.dx, nopt, p.delta_thresh, p.done, p.info);
    return;
  }

  // ---- back substitution x = L^-T y on CTA 0 -------------------------------
  // Thread o = k*32 + m keeps z_k[m] (register o / kCT); warp k % kWarps owns block k.
  double* xall = own;  // [Tc*32] (the tile storage is free now)
  const double* gt = p.gt;
  double z[kZ];
#pragma unroll
  for (int qq = 0; qq < kZ; ++qq) {
    z[qq] = 0.0;
    const int o = tid + qq * kCT, k = o >> 5, m = o & 31;
    if (k < Tc) {
      const int nvk = min(TB, n - TB * k);
      double v;
      if (k == T - 1) v = __ldcg(ytail + m);
      else v = __ldcg(gt + (size_t)((T - 1) * T / 2 + k) * kTile + swz(rl, m));
      z[qq] = (m < nvk) ? v : 0.0;
    }
  }
  const int lane = tid & 31, warp = tid >> 5;
  for (int i = Tc - 1; i >= 0; --i) {
    if (warp == (i % kWarps)) {
      const int nvi = min(TB, n - TB * i);
      double zi = 0.0;
#pragma unroll
      for (int qq = 0; qq < kZ; ++qq)
        if (qq == i / kWarps) zi = z[qq];
      s_z[lane] = zi;
      __syncwarp();
      const double* Wt = gt + (size_t)(i * (i + 1) / 2 + i) * kTile;
      double x = 0.0;
      for (int r = lane; r < TB; ++r) x = fma(__ldcg(Wt + swz(r, lane)), s_z[r], x);
      if (lane >= nvi) x = 0.0;
      s_x[lane] = x;
      xall[i * TB + lane] = x;
    }
    __syncthreads();
#pragma unroll
    for (int qq = 0; qq < kZ; ++qq) {
      const int o = tid + qq * kCT, k = o >> 5, m = o & 31;
      if (k < i) {
        const double* Lt = gt + (size_t)(i * (i + 1) / 2 + k) * kTile;
        double acc = 0.0;
#pragma unroll 8
        for (int r = 0; r < TB; ++r) acc = fma(__ldcg(Lt + swz(r, m)), s_x[r], acc);
        z[qq] -= acc;
      }
    }
    __syncthreads();
  }
  finish_update<kCT>(xall, p.Twc, p.dx, n, nopt, p.num_fix, p.delta_thresh, p.done, p.info);
}

}  // namespace

// Bytes of global scratch the cluster path needs for a system of dimension n.
size_t gn_chol_scratch_bytes(int n) {
  const int T = (n + 1 + TB - 1) / TB;
  return ((size_t)T * (T + 1) / 2 * kTile + TB) * sizeof(double);
}

namespace {

struct ClusterPlan {
  int C = 0;  // 0 = use the single-CTA global-memory kernel
  int max_own = 0, npanel = 0;
  size_t smem = 0;
};

// Chooses the cluster size for T tile rows (cached per T; the library serves one device
// per process).
ClusterPlan plan_cluster(int T, int Tc) {
  static ClusterPlan cache[40];
  static bool have[40] = {};
  static bool attr_set = false;
  ClusterPlan none;
  if (T >= 40 || Tc > 32) return none;
  if (have[T]) return cache[T];
  have[T] = true;
  cache[T] = none;
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, chol_cluster_kernel) != cudaSuccess) return none;
  const size_t limit = 232448 - fa.sharedSizeBytes;
  if (!attr_set) {
    if (cudaFuncSetAttribute(chol_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)limit) != cudaSuccess ||
        cudaFuncSetAttribute(chol_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed,
                             1) != cudaSuccess) {
      cudaGetLastError();
      return none;
    }
    attr_set = true;
  }
  const int ntiles = T * (T + 1) / 2;
  int C = ntiles >= 64 ? 8 : ntiles >= 24 ? 4 : ntiles >= 8 ? 2 : 1;
  for (; C <= 16; C *= 2) {
    ClusterPlan pl;
    pl.C = C;
    pl.max_own = (ntiles + C - 1) / C;
    pl.npanel = T - 1 > 2 ? T - 1 : 2;
    pl.smem = (size_t)(pl.max_own + pl.npanel + 1) * kTile * sizeof(double);
    if (pl.smem > limit || pl.max_own > kMaxOwn) continue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(C);
    cfg.blockDim = dim3(kCT);
    cfg.dynamicSmemBytes = pl.smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int nclusters = 0;
    if (cudaOccupancyMaxActiveClusters(&nclusters, chol_cluster_kernel, &cfg) != cudaSuccess ||
        nclusters < 1) {
      cudaGetLastError();
      continue;
    }
    cache[T] = pl;
    break;
  }
  return cache[T];
}

}  // namespace

int launch_gn_solve_retract(double* Hg, float* Twc, float* dx, int K, int num_fix,
                            float delta_thresh, int32_t* done, int32_t* info, double* chol_scratch,
                            cudaStream_t s) {
  const int n = 7 * (K - num_fix);
  if (n <= 0) return FRODO_OK;
  const int T = (n + 1 + TB - 1) / TB, Tc = (n + TB - 1) / TB;
  const ClusterPlan pl = chol_scratch != nullptr ? plan_cluster(T, Tc) : ClusterPlan();
  if (pl.C > 0) {
    ClusterArgs a;
    a.Hg = Hg; a.gt = chol_scratch; a.Twc = Twc; a.dx = dx; a.done = done; a.info = info;
    a.n = n; a.K = K; a.num_fix = num_fix; a.T = T; a.Tc = Tc;
    a.max_own = pl.max_own; a.npanel = pl.npanel; a.delta_thresh = delta_thresh;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pl.C);
    cfg.blockDim = dim3(kCT);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = pl.C;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    FRODO_CUDA_TRY(cudaLaunchKernelEx(&cfg, chol_cluster_kernel, a));
    return FRODO_OK;
  }
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
