// Keyframe-id indexing (get_unique_kf_idx + create_inds, gn_kernels.cu:203-212)
// and assembly of the dense fp64 normal equations from the per-edge blocks
// (SparseBlock::update_lhs / update_rhs, gn_kernels.cu:72-122).
#include "gn_internal.cuh"

namespace frodo {

namespace {

constexpr int kIdxThreads = 1024;

// Single CTA: bitonic sort of the 2E ids (padded to P = 2^m with INT64_MAX),
// prefix count of distinct values, then rank of every edge endpoint.
__global__ void __launch_bounds__(kIdxThreads)
edge_index_kernel(const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int E, int P,
                  int64_t* __restrict__ keys_g, int32_t* __restrict__ pref_g,
                  int32_t* __restrict__ ii_row, int32_t* __restrict__ jj_row,
                  int32_t* __restrict__ n_unique, int use_smem,
                  const int64_t* __restrict__ q_ids, int nq, int32_t* __restrict__ q_rows) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int64_t* keys = use_smem ? reinterpret_cast<int64_t*>(smem_raw) : keys_g;
  int32_t* pref = use_smem ? reinterpret_cast<int32_t*>(smem_raw + (size_t)P * 8) : pref_g;
  __shared__ int32_t scan[kIdxThreads];
  const int tid = threadIdx.x;
  const int n2 = 2 * E;
  for (int k = tid; k < P; k += kIdxThreads)
    keys[k] = k < E ? ii[k] : (k < n2 ? jj[k - E] : INT64_MAX);
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int k = tid; k < P / 2; k += kIdxThreads) {
        const int lo = 2 * k - (k & (stride - 1));
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const int64_t a = keys[lo], b = keys[hi];
        if ((a > b) == up) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  }
  // head flags -> exclusive prefix: pref[k] = number of distinct values among sorted[0..k)
  const int per = (P + kIdxThreads - 1) / kIdxThreads;
  const int beg = tid * per, end = min(P, beg + per);
  int cnt = 0;
  for (int k = beg; k < end; ++k)
    if (k < n2 && (k == 0 || keys[k] != keys[k - 1])) ++cnt;
  scan[tid] = cnt;
  __syncthreads();
  for (int off = 1; off < kIdxThreads; off <<= 1) {
    const int x = tid >= off ? scan[tid - off] : 0;
    __syncthreads();
    scan[tid] += x;
    __syncthreads();
  }
  int run = scan[tid] - cnt;
  for (int k = beg; k < end; ++k) {
    pref[k] = run;
    if (k < n2 && (k == 0 || keys[k] != keys[k - 1])) ++run;
  }
  if (tid == kIdxThreads - 1 && n_unique != nullptr) *n_unique = scan[tid];
  __syncthreads();
  for (int k = tid; k < n2; k += kIdxThreads) {
    const int64_t key = k < E ? ii[k] : jj[k - E];
    int lo = 0, hi = n2;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (keys[mid] < key) lo = mid + 1; else hi = mid;
    }
    const int r = pref[lo];
    if (k < E) ii_row[k] = r; else jj_row[k - E] = r;
  }
  // extra queries (odometry edge endpoints): torch.searchsorted(unique, id), also for ids
  // that are not keyframes of any visual edge (gn_kernels.cu:1817-1829)
  const int total = scan[kIdxThreads - 1];
  for (int k = tid; k < nq; k += kIdxThreads) {
    const int64_t key = q_ids[k];
    int lo = 0, hi = n2;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (keys[mid] < key) lo = mid + 1; else hi = mid;
    }
    q_rows[k] = lo < n2 ? pref[lo] : total;
  }
}

// One CTA per lower block (a >= b) of H; also writes the mirrored block, and the
// diagonal CTAs write g.  Edges are scanned in index order, so the fp64 sums are
// order-deterministic.  H = sum_e (e_i - e_j)(e_i - e_j)^T (x) A_e,
// g = sum_e (e_j - e_i) (x) gj_e  -- the closed form of update_lhs/update_rhs with
// H_ii = H_jj = A, H_ij = H_ji = -A, g_i = -g_j.
__global__ void __launch_bounds__(64)
assemble_kernel(const double* __restrict__ edge_blocks, const int32_t* __restrict__ ii_row,
                const int32_t* __restrict__ jj_row, int E, int nblk, int num_fix,
                double* __restrict__ Hg, const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  // decode lower-triangular block index
  const int bid = blockIdx.x;
  int a = (int)((sqrtf(8.0f * (float)bid + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= bid) ++a;
  while (a * (a + 1) / 2 > bid) --a;
  const int b = bid - a * (a + 1) / 2;
  const int n = 7 * nblk;
  const int tid = threadIdx.x;
  const int r = tid / 7, c = tid % 7;
  double acc = 0.0;
  int sidx = 0;
  if (tid < 49) {
    const int lo = min(r, c), hi = max(r, c);
    sidx = lo * 7 - lo * (lo - 1) / 2 + (hi - lo);
  } else if (tid < 56) {
    sidx = 28 + (tid - 49);
  }
  const bool is_g = (tid >= 49 && tid < 56);
  // The 64 threads classify 64 edges at a time; ballots give the (few) edges that touch
  // this block, visited in edge order by every accumulating thread.
  __shared__ signed char s_ch[64], s_cg[64];
  __shared__ unsigned s_mask[2];
  for (int base = 0; base < E; base += 64) {
    const int e = base + tid;
    int ch = 0, cg = 0;
    if (e < E) {
      const int io = ii_row[e] - num_fix, jo = jj_row[e] - num_fix;
      ch = (io == a && io == b) + (jo == a && jo == b) - (io == a && jo == b) -
           (jo == a && io == b);
      if (a == b) cg = (jo == a) - (io == a);
    }
    s_ch[tid] = (signed char)ch;
    s_cg[tid] = (signed char)cg;
    const unsigned m = __ballot_sync(0xffffffffu, ch != 0 || cg != 0);
    if ((tid & 31) == 0) s_mask[tid >> 5] = m;
    __syncthreads();
    if (tid < 56) {
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        unsigned mm = s_mask[w];
        while (mm) {
          const int k = w * 32 + __ffs(mm) - 1;
          mm &= mm - 1;
          const int coef = is_g ? s_cg[k] : s_ch[k];
          if (coef != 0)
            acc += (double)coef * edge_blocks[(size_t)(base + k) * kGnBlock + sidx];
        }
      }
    }
    __syncthreads();
  }
  if (tid < 49) {
    Hg[(size_t)(7 * a + r) * n + (7 * b + c)] = acc;
    if (a != b) Hg[(size_t)(7 * b + c) * n + (7 * a + r)] = acc;
  } else if (is_g && a == b) {
    Hg[(size_t)n * n + 7 * a + (tid - 49)] = acc;
  }
}

}  // namespace

int launch_gn_edge_index(const int64_t* ii, const int64_t* jj, int E, int32_t* ii_row,
                         int32_t* jj_row, int32_t* n_unique, cudaStream_t s,
                         const int64_t* q_ids, int nq, int32_t* q_rows) {
  if (E == 0) {
    if (n_unique) FRODO_CUDA_TRY(cudaMemsetAsync(n_unique, 0, sizeof(int32_t), s));
    if (nq > 0) FRODO_CUDA_TRY(cudaMemsetAsync(q_rows, 0, (size_t)nq * sizeof(int32_t), s));
    return FRODO_OK;
  }
  int P = 2;
  while (P < 2 * E) P <<= 1;
  const size_t need = (size_t)P * 12;
  if (need > 200 * 1024) {
    set_error("edge index: more than %d edges are not supported", (200 * 1024 / 12) / 2);
    return FRODO_ERR_ARG;
  }
  static PerDeviceOnce once;
  int dev;
  if (once.needed(&dev)) {
    FRODO_CUDA_TRY(cudaFuncSetAttribute(edge_index_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    once.mark(dev);
  }
  edge_index_kernel<<<1, kIdxThreads, need, s>>>(ii, jj, E, P, nullptr, nullptr, ii_row, jj_row,
                                                  n_unique, 1, q_ids, nq, q_rows);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_gn_assemble(const double* edge_blocks, const int32_t* ii_row, const int32_t* jj_row,
                       int E, int K, int num_fix, double* Hg, const int32_t* done,
                       cudaStream_t s) {
  const int nblk = K - num_fix;
  if (nblk <= 0) return FRODO_OK;
  const int nb = nblk * (nblk + 1) / 2;
  assemble_kernel<<<nb, 64, 0, s>>>(edge_blocks, ii_row, jj_row, E, nblk, num_fix, Hg, done);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
