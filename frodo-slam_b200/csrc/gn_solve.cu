// Dense fp64 Cholesky solve of the assembled normal equations, pose retraction
// and the convergence test, in one kernel (no host round trip):
//   SparseBlock::solve (gn_kernels.cu:163-199; failure -> dx = 0),
//   dx = -x -> fp32 (:977, :191-192), pose_retr_kernel (:522-560),
//   ||dx||_2 < delta_thresh (:986-989).
// Hg is the (n+1) x n row-major matrix [H ; g^T]: factorising it as if g were
// one more matrix row yields y = L^-1 g in the last row for free.
#include "gn_internal.cuh"
#include "gn_pose_update.cuh"

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

using namespace pose_update;

// Single CTA, right-looking blocked Cholesky on the (n+1) x n matrix in global
// memory (L2 resident), lower triangle only.
__global__ void __launch_bounds__(kSolveThreads)
solve_retract_kernel(double* __restrict__ A, float* __restrict__ Twc, float* __restrict__ dx,
                     int n, int K, int num_fix, float delta_thresh, int32_t* __restrict__ done,
                     int32_t* __restrict__ info, const int32_t* __restrict__ pose_rows) {
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
        if (!(d > 0.0) || !isfinite(d)) {
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

  finish_update<kSolveThreads>(y, Twc, dx, n, nopt, num_fix, delta_thresh, done, info, pose_rows);
}


// ---------------------------------------------------------------------------------------
// Small systems (n <= 63: up to 9 free keyframes, i.e. the sliding-window local optimisation of
// the fork's tracker, tracker.py:27-60): the multi-CTA machinery of gn_chol_flow.cu (cooperative
// launch, L2 tile array, flags) is pointless for two 32x32 tiles.  Here the whole
// (n+1) x n system [H ; g^T] sits in the shared memory of one CTA (row pitch 65 doubles:
// column accesses are conflict-free) and a right-looking factorisation runs with two block
// barriers per pivot: every thread forms 1/sqrt(a_jj) redundantly, 64 threads scale column j,
// a 16 x 16 thread grid applies the rank-1 update.  Rolled loops on purpose: a fully unrolled
// register-resident variant (row t in the registers of thread t) was 2x SLOWER than the cluster
// path -- 300 KB of straight-line code executed once is bound by instruction fetch.
// The rhs rides along as row n (y = L^-1 g); the back substitution is column-oriented.
constexpr int kSmallN = 63;
constexpr int kSmallThreads = 256;
constexpr int kSmallLd = 65;

__global__ void __launch_bounds__(kSmallThreads)
solve_small_kernel(const double* __restrict__ A, float* __restrict__ Twc, float* __restrict__ dx,
                   int n, int K, int num_fix, float delta_thresh, int32_t* __restrict__ done,
                   int32_t* __restrict__ info, const int32_t* __restrict__ pose_rows) {
  if (done != nullptr && *done != 0) return;
  __shared__ double M[(kSmallN + 1) * kSmallLd];  // rows 0..n, lower triangle used
  __shared__ double lcol[kSmallN + 1];            // column j of L during step j
  __shared__ double s_inv[kSmallN + 1];
  __shared__ double xs[kSmallN + 1];
  const int t = threadIdx.x;
  const int rows = n + 1;
  // coalesced load, 4 independent loads in flight per thread
  for (int r0 = 0; r0 < rows; r0 += 4 * (kSmallThreads / 64)) {
    double v[4];
    const int c = t & 63, rr = t >> 6;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + rr + u * (kSmallThreads / 64);
      v[u] = (r < rows && c < n) ? A[(size_t)r * n + c] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + rr + u * (kSmallThreads / 64);
      if (r < rows && c < n) M[r * kSmallLd + c] = v[u];
    }
  }
  __syncthreads();
  bool fail = false;
  const int ty = t >> 4, tx = t & 15;
  for (int j = 0; j < n; ++j) {
    const double d = M[j * kSmallLd + j];
    fail = !(d > 0.0) || !isfinite(d);
    if (fail) break;  // uniform: every thread read the same pivot
    const double inv = rsqrt(d);
    if (t >= j && t < rows) {
      const double l = (t == j) ? d * inv : M[t * kSmallLd + j] * inv;
      lcol[t] = l;
      if (t == j) s_inv[j] = inv;
    }
    __syncthreads();
    if (t >= j && t < rows) M[t * kSmallLd + j] = lcol[t];
    for (int r = j + 1 + ty; r < rows; r += 16) {
      const double lr = lcol[r];
      for (int c = j + 1 + tx; c <= r && c < n; c += 16)
        M[r * kSmallLd + c] = fma(-lr, lcol[c], M[r * kSmallLd + c]);
    }
    __syncthreads();
  }
  const int nopt = K - num_fix;
  if (fail) {
    fail_update<kSmallThreads>(dx, nopt, delta_thresh, done, info);
    return;
  }
  // x = L^-T y with y = row n; thread i < n keeps y_i, x_j is broadcast through xs
  double yi = t < n ? M[n * kSmallLd + t] : 0.0;
  for (int j = n - 1; j >= 0; --j) {
    if (t == j) xs[j] = yi * s_inv[j];
    __syncthreads();
    if (t < j) yi = fma(-M[j * kSmallLd + t], xs[j], yi);
  }
  __syncthreads();
  finish_update<kSmallThreads>(xs, Twc, dx, n, nopt, num_fix, delta_thresh, done, info, pose_rows);
}

}  // namespace

int launch_gn_solve_retract(const GnSystem& sys, float* Twc, float* dx, int K, int num_fix,
                            float delta_thresh, int32_t* done, int32_t* info, double* chol_scratch,
                            int n_max, cudaStream_t s, const int32_t* pose_rows) {
  const int n = 7 * (K - num_fix);
  if (n <= 0) return FRODO_OK;
  double* Hg = sys.Hg;
  auto need_dense = [&]() -> int {
    if (sys.dense) return FRODO_OK;
    if (sys.wait_counter != nullptr) {
      const int wrc = launch_gn_peer_wait(sys.wait_counter, sys.wait_target, done, info, s);
      if (wrc != FRODO_OK) return wrc;
    }
    return launch_gn_assemble(sys.edge_blocks, sys.ii_row, sys.jj_row, sys.E, K, num_fix, Hg, done,
                              s);
  };
  if (n <= kSmallN) {  // sliding-window sized systems: one CTA, system in shared memory
    if (!sys.dense) {  // straight from the per-edge blocks (gn_chol_flow.cu, chol_small_kernel)
      if (sys.wait_counter != nullptr) {
        const int wrc = launch_gn_peer_wait(sys.wait_counter, sys.wait_target, done, info, s);
        if (wrc != FRODO_OK) return wrc;
      }
      int handled = 0;
      const int src = launch_chol_small(sys.edge_blocks, sys.ii_row, sys.jj_row, sys.E, Twc, dx, K,
                                        num_fix, delta_thresh, done, info, &handled, s, pose_rows);
      if (src != FRODO_OK || handled) return src;
    }
    const int rc = need_dense();
    if (rc != FRODO_OK) return rc;
    solve_small_kernel<<<1, kSmallThreads, 0, s>>>(Hg, Twc, dx, n, K, num_fix, delta_thresh, done,
                                                   info, pose_rows);
    FRODO_CUDA_TRY(cudaGetLastError());
    return FRODO_OK;
  }
  // fast path: dataflow tiled Cholesky over the whole GPU (gn_chol_flow.cu), tiles built from
  // the dense system or straight from the per-edge blocks
  int handled = 0;
  int rc = launch_chol_flow(sys.dense ? Hg : nullptr, sys.edge_blocks, sys.ii_row, sys.jj_row,
                            sys.E, Twc, dx, K, num_fix, delta_thresh, done, info, chol_scratch,
                            n_max, &handled, s, pose_rows, sys.dense ? nullptr : sys.wait_counter,
                            sys.wait_target);
  if (rc != FRODO_OK || handled) return rc;
  // fallback: no cooperative launch on this device, no scratch, or beyond the tile plan
  rc = need_dense();
  if (rc != FRODO_OK) return rc;
  static PerDeviceOnce once;
  int dev;
  if (once.needed(&dev)) {
    FRODO_CUDA_TRY(cudaFuncSetAttribute(solve_retract_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(SolveSmem)));
    once.mark(dev);
  }
  solve_retract_kernel<<<1, kSolveThreads, sizeof(SolveSmem), s>>>(Hg, Twc, dx, n, K, num_fix,
                                                                  delta_thresh, done, info,
                                                                  pose_rows);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
