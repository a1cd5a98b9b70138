// Gauss-Newton accumulation: per directed edge (i,j) the Huber-weighted
// residual Jacobians of ray_align_kernel / point_align_kernel /
// calib_proj_kernel (gn_kernels.cu:576-907, 1936-2204, 2295-2607) reduced to
// one 7x7 block and one 7-vector per edge.
//
// Two exact identities of the reference math make the kernel stream-bound:
//  (1) Ji = -Jj (gn_kernels.cu:769), so H_ii = H_jj = -H_ij = -H_ji =: A and
//      g_i = -g_j: one symmetric 7x7 (28 values) + 7 values per edge suffice;
//  (2) Jj = M x with M = apply_Sim3_adj_inv(T_i) linear and constant per edge
//      (gn_kernels.cu:319-339), so A = M (sum w x x^T) M^T and g_j = M sum(w e x):
//      threads accumulate the sparse local rows x and M is applied once per edge.
// Sums over points: fp32 per thread (<= 32 points each) -> fp32 warp/CTA tree
// -> fp64 across CTAs; M S M^T in fp64.
#include "common.cuh"
#include "gn_internal.cuh"
#include "ptx_async.cuh"

namespace frodo {

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

struct Sim3f {
  float t[3], q[4], s;
};

__device__ __forceinline__ Sim3f load_pose(const float* __restrict__ Twc, int row) {
  Sim3f T;
  const float4 a = __ldg(reinterpret_cast<const float4*>(Twc + (size_t)row * 8));
  const float4 b = __ldg(reinterpret_cast<const float4*>(Twc + (size_t)row * 8) + 1);
  T.t[0] = a.x; T.t[1] = a.y; T.t[2] = a.z;
  T.q[0] = a.w; T.q[1] = b.x; T.q[2] = b.y; T.q[3] = b.z;
  T.s = b.w;
  return T;
}

// actSO3, gn_kernels.cu:237-247
__device__ __forceinline__ void act_so3(const float* q, const float* X, float* Y) {
  const float uv0 = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  const float uv1 = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  const float uv2 = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  Y[1] = X[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  Y[2] = X[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
}

// relSim3, gn_kernels.cu:294-314: T_ij = T_i^-1 T_j
__device__ __forceinline__ Sim3f rel_sim3(const Sim3f& Ti, const Sim3f& Tj) {
  Sim3f R;
  const float si_inv = 1.0f / Ti.s;
  R.s = si_inv * Tj.s;
  const float qi[4] = {-Ti.q[0], -Ti.q[1], -Ti.q[2], Ti.q[3]};
  const float* qj = Tj.q;
  R.q[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  R.q[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  R.q[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  R.q[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
  const float d[3] = {Tj.t[0] - Ti.t[0], Tj.t[1] - Ti.t[1], Tj.t[2] - Ti.t[2]};
  act_so3(qi, d, R.t);
  R.t[0] *= si_inv; R.t[1] *= si_inv; R.t[2] *= si_inv;
  return R;
}

// huber(), gn_kernels.cu:214-217 (k = 1.345)
__device__ __forceinline__ float huber(float r) {
  const float a = fabsf(r);
  return a < 1.345f ? 1.0f : __fdividef(1.345f, a);
}

// S (28 upper-triangular entries, row-major) += w x x^T and v += (w e) x, using
// only the entries of x named in MASK (the others are structurally zero).
template <unsigned MASK>
__device__ __forceinline__ void accum_row(float (&S)[28], float (&v)[7], const float (&x)[7],
                                          float w, float e) {
  const float we = w * e;
  float wx[7];
#pragma unroll
  for (int n = 0; n < 7; ++n)
    if ((MASK >> n) & 1u) wx[n] = w * x[n];
  int l = 0;
#pragma unroll
  for (int n = 0; n < 7; ++n) {
#pragma unroll
    for (int m = n; m < 7; ++m) {
      if (((MASK >> n) & 1u) && ((MASK >> m) & 1u)) S[l] = fmaf(wx[n], x[m], S[l]);
      ++l;
    }
  }
#pragma unroll
  for (int n = 0; n < 7; ++n)
    if ((MASK >> n) & 1u) v[n] = fmaf(we, x[n], v[n]);
}

struct EdgeConst {
  Sim3f Tij;
  float sa_inv, sb_inv, C_thresh, Q_thresh;
  float fx, fy, cx, cy;
  float ulo, uhi, vlo, vhi, z_eps;
  int width;
};

template <int MODE>
__device__ __forceinline__ void point_terms(const EdgeConst& c, float (&S)[28], float (&v)[7],
                                            const float* Xi, const float* Xj, float q, float ci,
                                            float cj, bool vm, long long ind) {
  float P[3];
  act_so3(c.Tij.q, Xj, P);
  P[0] = fmaf(P[0], c.Tij.s, c.Tij.t[0]);
  P[1] = fmaf(P[1], c.Tij.s, c.Tij.t[1]);
  P[2] = fmaf(P[2], c.Tij.s, c.Tij.t[2]);
  bool valid = vm & (q > c.Q_thresh) & (ci > c.C_thresh) & (cj > c.C_thresh);
  const float sq = sqrtf(q);
  float x[7];
  if (MODE == FRODO_GN_RAYS) {
    const float n2i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
    const float ni_inv = rsqrtf(n2i);
    const float ni = n2i * ni_inv;
    const float n2j = P[0] * P[0] + P[1] * P[1] + P[2] * P[2];
    const float nj_inv = rsqrtf(n2j);
    const float nj = n2j * nj_inv;
    const float r0 = P[0] * nj_inv, r1 = P[1] * nj_inv, r2 = P[2] * nj_inv;
    const float e0 = fmaf(-Xi[0], ni_inv, r0);
    const float e1 = fmaf(-Xi[1], ni_inv, r1);
    const float e2 = fmaf(-Xi[2], ni_inv, r2);
    const float e3 = nj - ni;
    const float sw_r = valid ? c.sa_inv * sq : 0.0f;
    const float sw_d = valid ? c.sb_inv * sq : 0.0f;
    const float wr = sw_r * sw_r, wd = sw_d * sw_d;
    const float w0 = huber(sw_r * e0) * wr;
    const float w1 = huber(sw_r * e1) * wr;
    const float w2 = huber(sw_r * e2) * wr;
    const float w3 = huber(sw_d * e3) * wd;
    // d r / d P = (I - r r^T) / |P|   (gn_kernels.cu:751-757)
    const float dxx = fmaf(-r0, r0, 1.0f) * nj_inv;
    const float dyy = fmaf(-r1, r1, 1.0f) * nj_inv;
    const float dzz = fmaf(-r2, r2, 1.0f) * nj_inv;
    const float dxy = -r0 * r1 * nj_inv;
    const float dxz = -r0 * r2 * nj_inv;
    const float dyz = -r1 * r2 * nj_inv;
    x[0] = dxx; x[1] = dxy; x[2] = dxz; x[3] = 0.f; x[4] = r2; x[5] = -r1; x[6] = 0.f;
    accum_row<0x37u>(S, v, x, w0, e0);  // {0,1,2,4,5}
    x[0] = dxy; x[1] = dyy; x[2] = dyz; x[3] = -r2; x[4] = 0.f; x[5] = r0; x[6] = 0.f;
    accum_row<0x2Fu>(S, v, x, w1, e1);  // {0,1,2,3,5}
    x[0] = dxz; x[1] = dyz; x[2] = dzz; x[3] = r1; x[4] = -r0; x[5] = 0.f; x[6] = 0.f;
    accum_row<0x1Fu>(S, v, x, w2, e2);  // {0,1,2,3,4}
    x[0] = r0; x[1] = r1; x[2] = r2; x[3] = 0.f; x[4] = 0.f; x[5] = 0.f; x[6] = nj;
    accum_row<0x47u>(S, v, x, w3, e3);  // {0,1,2,6}
  } else if (MODE == FRODO_GN_POINTS) {
    const float e0 = P[0] - Xi[0], e1 = P[1] - Xi[1], e2 = P[2] - Xi[2];
    const float sw = valid ? c.sa_inv * sq : 0.0f;
    const float ws = sw * sw;
    const float w0 = huber(sw * e0) * ws;
    const float w1 = huber(sw * e1) * ws;
    const float w2 = huber(sw * e2) * ws;
    x[0] = 1.f; x[1] = 0.f; x[2] = 0.f; x[3] = 0.f; x[4] = P[2]; x[5] = -P[1]; x[6] = P[0];
    accum_row<0x71u>(S, v, x, w0, e0);  // {0,4,5,6}
    x[0] = 0.f; x[1] = 1.f; x[2] = 0.f; x[3] = -P[2]; x[4] = 0.f; x[5] = P[0]; x[6] = P[1];
    accum_row<0x6Au>(S, v, x, w1, e1);  // {1,3,5,6}
    x[0] = 0.f; x[1] = 0.f; x[2] = 1.f; x[3] = P[1]; x[4] = -P[0]; x[5] = 0.f; x[6] = P[2];
    accum_row<0x5Cu>(S, v, x, w2, e2);  // {2,3,4,6}
  } else {  // FRODO_GN_CALIB, gn_kernels.cu:2424-2559
    const int u_t = (int)(ind % c.width);
    const int v_t = (int)(ind / c.width);
    const bool valid_z = (P[2] > c.z_eps) && (Xi[2] > c.z_eps);
    const float zinv = valid_z ? 1.0f / P[2] : 0.0f;
    const float zj_log = valid_z ? logf(P[2]) : 0.0f;
    const float zi_log = valid_z ? logf(Xi[2]) : 0.0f;
    const float xz = P[0] * zinv, yz = P[1] * zinv;
    const float u = fmaf(c.fx, xz, c.cx);
    const float vv = fmaf(c.fy, yz, c.cy);
    const bool valid_u = (u > c.ulo) && (u < c.uhi);
    const bool valid_v = (vv > c.vlo) && (vv < c.vhi);
    const float e0 = u - (float)u_t, e1 = vv - (float)v_t, e2 = zj_log - zi_log;
    valid = valid & valid_u & valid_v & valid_z;
    const float sw_p = valid ? c.sa_inv * sq : 0.0f;
    const float sw_d = valid ? c.sb_inv * sq : 0.0f;
    const float wp = sw_p * sw_p, wd = sw_d * sw_d;
    const float w0 = huber(sw_p * e0) * wp;
    const float w1 = huber(sw_p * e1) * wp;
    const float w2 = huber(sw_d * e2) * wd;
    x[0] = c.fx * zinv; x[1] = 0.f; x[2] = -c.fx * xz * zinv; x[3] = -c.fx * xz * yz;
    x[4] = c.fx * fmaf(xz, xz, 1.0f); x[5] = -c.fx * yz; x[6] = 0.f;
    accum_row<0x3Du>(S, v, x, w0, e0);  // {0,2,3,4,5}
    x[0] = 0.f; x[1] = c.fy * zinv; x[2] = -c.fy * yz * zinv; x[3] = -c.fy * fmaf(yz, yz, 1.0f);
    x[4] = c.fy * xz * yz; x[5] = c.fy * xz; x[6] = 0.f;
    accum_row<0x3Eu>(S, v, x, w1, e1);  // {1,2,3,4,5}
    x[0] = 0.f; x[1] = 0.f; x[2] = zinv; x[3] = yz; x[4] = -xz; x[5] = 0.f; x[6] = 1.f;
    accum_row<0x5Cu>(S, v, x, w2, e2);  // {2,3,4,6}
  }
}

// ---------------------------------------------------------------------------
// rays, fast path helpers (the kernel and its arithmetic are documented further down).
// Weights: w = huber(|e| sqrt(q)/sigma) q/sigma^2 = min(1, 1.345 rsqrt(t)) wq with
// t = wq e^2, wq = q/sigma^2 (same value as gn_kernels.cu:214-217,730-741, no sqrt(q)).
// ---------------------------------------------------------------------------
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 bc(float a) { return make_float2(a, a); }

__device__ __forceinline__ float rsqrt_fast(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct RaysConst {  // per-edge constants of the packed kernels (all three modes)
  float M[9];  // s_ij R(q_ij)
  float t[3];
  float sa2, sb2, C_thresh, Q_thresh;
  float fx, fy, cx, cy, ulo, uhi, vlo, vhi, z_eps;  // calib only
  unsigned width, width_rcp;                         // calib only; floor(2^32 / width)
};

__host__ __device__ constexpr int tri(int n, int m) { return n * 7 - n * (n - 1) / 2 + (m - n); }

// Adjoint map M of pose i (7x7, fp64):  Jj = M x  (apply_Sim3_adj_inv, :319-339)
__device__ void adjoint_matrix(const Sim3f& Ti, double* M /*[49]*/) {
  const double qx = Ti.q[0], qy = Ti.q[1], qz = Ti.q[2], qw = Ti.q[3];
  double R[9];
  // R = I + 2 w [q]x + 2 [q]x^2  (same map as actSO3, without normalising q)
  R[0] = 1 - 2 * (qy * qy + qz * qz); R[1] = 2 * (qx * qy - qz * qw); R[2] = 2 * (qx * qz + qy * qw);
  R[3] = 2 * (qx * qy + qz * qw); R[4] = 1 - 2 * (qx * qx + qz * qz); R[5] = 2 * (qy * qz - qx * qw);
  R[6] = 2 * (qx * qz - qy * qw); R[7] = 2 * (qy * qz + qx * qw); R[8] = 1 - 2 * (qx * qx + qy * qy);
  const double s_inv = 1.0 / (double)Ti.s;
  const double t[3] = {Ti.t[0], Ti.t[1], Ti.t[2]};
  for (int k = 0; k < 49; ++k) M[k] = 0.0;
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      M[r * 7 + c] = s_inv * R[r * 3 + c];
      M[(3 + r) * 7 + 3 + c] = R[r * 3 + c];
    }
  for (int c = 0; c < 3; ++c) {
    const double Rc[3] = {R[0 + c], R[3 + c], R[6 + c]};  // column c of R
    M[3 * 7 + c] = s_inv * (t[1] * Rc[2] - t[2] * Rc[1]);
    M[4 * 7 + c] = s_inv * (t[2] * Rc[0] - t[0] * Rc[2]);
    M[5 * 7 + c] = s_inv * (t[0] * Rc[1] - t[1] * Rc[0]);
    M[6 * 7 + c] = s_inv * (t[0] * Rc[0] + t[1] * Rc[1] + t[2] * Rc[2]);
  }
  M[6 * 7 + 6] = 1.0;
}

// CTA reduction of the 35 per-thread sums (warp shuffles, one smem stage), fp32 partial
// per chunk; the last CTA of the edge (ticket) sums the chunks in fixed order in fp64 and
// applies the adjoint: A = M S M^T, g = M v.
// Slot (e, chunk) of `partials` is at (e * stride + chunk); stride >= nchunks.
// BAR = 0: the whole CTA (THREADS == blockDim.x); otherwise named barrier BAR over the first
// THREADS threads (the consumer warps of a warp-specialised CTA).
template <int THREADS, int BAR>
__device__ __forceinline__ void epi_sync() {
  if constexpr (BAR == 0) __syncthreads();
  else ptx::named_bar_sync<BAR, THREADS>();
}

// One arrival of this rank on rank r's counter (release: ordered after this CTA's block stores).
__device__ __forceinline__ void peer_arrive(const GnPeerOut& peer, int r) {
  uint32_t* cnt = reinterpret_cast<uint32_t*>(static_cast<char*>(peer.bufs[r]) + peer.cnt_offset);
  asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
}
// A skipped pass (convergence flag set) still arrives once per edge, so that the cumulative
// arrival counts stay E per epoch on every rank.
__device__ __forceinline__ void peer_skip(const GnPeerOut& peer) {
  if (peer.world > 0 && blockIdx.x == 0 && threadIdx.x < peer.world) peer_arrive(peer, threadIdx.x);
}

template <int THREADS, int BAR = 0>
__device__ __forceinline__ void edge_epilogue(float (&S)[28], float (&v)[7], const Sim3f& Ti,
                                              int e, int chunk, int nchunks, int stride,
                                              float* __restrict__ partials,
                                              int32_t* __restrict__ tickets,
                                              double* __restrict__ edge_blocks,
                                              const GnPeerOut& peer) {
  __shared__ float warp_part[THREADS / 32][kGnBlock];
  __shared__ int is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {
    // entries 0..31 (28 of S, v[0..3]): reduce-scatter, lane m ends with the warp total of
    // entry m (31 exchanges); v[4..6]: butterflies.
    float a[32];
#pragma unroll
    for (int l = 0; l < 28; ++l) a[l] = S[l];
#pragma unroll
    for (int l = 0; l < 4; ++l) a[28 + l] = v[l];
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
      const bool up = (lane & half) != 0;
#pragma unroll
      for (int i = 0; i < half; ++i) {
        const float send = up ? a[i] : a[i + half];
        const float keep = up ? a[i + half] : a[i];
        a[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
      }
    }
    warp_part[warp][lane] = a[0];
#pragma unroll
    for (int l = 4; l < 7; ++l) {
      float x = v[l];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == 0) warp_part[warp][28 + l] = x;
    }
  }
  epi_sync<THREADS, BAR>();
  float* __restrict__ my_part = partials + ((size_t)e * stride + chunk) * kGnBlock;
  if (threadIdx.x < kGnBlock) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) x += warp_part[w][threadIdx.x];
    my_part[threadIdx.x] = x;
  }
  __threadfence();
  epi_sync<THREADS, BAR>();
  if (threadIdx.x == 0) {
    const int t = atomicAdd(tickets + e, 1);
    is_last = (t == nchunks - 1);
  }
  epi_sync<THREADS, BAR>();
  if (!is_last) return;
  __threadfence();

  // Last CTA of this edge: fixed-order fp64 sum over chunks, then A = M S M^T, g = M v.
  __shared__ double Sd[kGnBlock];
  __shared__ double Sfull[49];
  __shared__ double Md[49];
  __shared__ double MS[49];
  if (threadIdx.x < kGnBlock) {
    const float* p = partials + (size_t)e * stride * kGnBlock + threadIdx.x;
    double acc = 0.0;
    for (int ch = 0; ch < nchunks; ++ch) acc += (double)__ldcg(p + (size_t)ch * kGnBlock);
    Sd[threadIdx.x] = acc;
  }
  if (threadIdx.x == (THREADS > 64 ? 64 : THREADS - 1)) adjoint_matrix(Ti, Md);
  if (threadIdx.x == 0) tickets[e] = 0;  // ready for the next iteration
  epi_sync<THREADS, BAR>();
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    const int a = min(r, cc), b = max(r, cc);
    Sfull[threadIdx.x] = Sd[a * 7 - a * (a - 1) / 2 + (b - a)];
  }
  epi_sync<THREADS, BAR>();
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    double acc = 0.0;
    for (int a = 0; a < 7; ++a) acc += Md[r * 7 + a] * Sfull[a * 7 + cc];
    MS[threadIdx.x] = acc;
  }
  epi_sync<THREADS, BAR>();
  int slot = -1;
  double acc = 0.0;
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    if (cc >= r) {
      for (int b = 0; b < 7; ++b) acc += MS[r * 7 + b] * Md[cc * 7 + b];
      slot = r * 7 - r * (r - 1) / 2 + (cc - r);
    }
  } else if (threadIdx.x < 56) {
    const int r = threadIdx.x - 49;
    for (int a = 0; a < 7; ++a) acc += Md[r * 7 + a] * Sd[28 + a];
    slot = 28 + r;
  }
  if (peer.world <= 0) {
    if (slot >= 0) edge_blocks[(size_t)e * kGnBlock + slot] = acc;
    return;
  }
  // Multi-GPU: the exchange step is fused here.  The edge's 35 values go straight
  // into EVERY rank's copy of the global per-edge block array over NVLink peer memory, then
  // one release-add per rank on its arrival counter; the consumer (the tile build of the
  // Cholesky, gn_chol_flow.cu) waits for the counter and assembles in fixed edge order, so all
  // ranks hold bit-identical systems.  280 bytes per edge per rank instead of an all-reduce of
  // the dense [H | g].
  if (slot >= 0) {
    const size_t at = peer.eb_offset + ((size_t)(peer.edge_base + e) * kGnBlock + slot) * sizeof(double);
    for (int r = 0; r < peer.world; ++r)
      *reinterpret_cast<double*>(static_cast<char*>(peer.bufs[r]) + at) = acc;
    __threadfence_system();
  }
  epi_sync<THREADS, BAR>();
  if (threadIdx.x < peer.world) peer_arrive(peer, threadIdx.x);
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 2)
gn_accumulate_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                     const float* __restrict__ Xs, const float* __restrict__ Cs,
                     const float* __restrict__ Kmat, const int32_t* __restrict__ ii_row,
                     const int32_t* __restrict__ jj_row, const int64_t* __restrict__ idx,
                     const uint8_t* __restrict__ valid_match, const float* __restrict__ Q, int N,
                     int pts_per_cta, float* __restrict__ partials, int32_t* __restrict__ tickets,
                     double* __restrict__ edge_blocks, const int32_t* __restrict__ done,
                     const GnPeerOut peer) {
  if (done != nullptr && *done != 0) {
    peer_skip(peer);
    return;
  }
  const int e = blockIdx.y;
  const int chunk = blockIdx.x;
  const int nchunks = gridDim.x;
  const int ix = ii_row[e], jx = jj_row[e];
  const Sim3f Ti = load_pose(Twc, ix);
  const Sim3f Tj = load_pose(Twc, jx);

  EdgeConst c;
  c.Tij = rel_sim3(Ti, Tj);
  c.sa_inv = 1.0f / prm.sigma_a;
  c.sb_inv = (MODE == FRODO_GN_POINTS) ? 0.0f : 1.0f / prm.sigma_b;
  c.C_thresh = prm.C_thresh;
  c.Q_thresh = prm.Q_thresh;
  c.width = prm.width > 0 ? prm.width : 1;
  if (MODE == FRODO_GN_CALIB) {
    c.fx = __ldg(Kmat + 0); c.fy = __ldg(Kmat + 4); c.cx = __ldg(Kmat + 2); c.cy = __ldg(Kmat + 5);
    c.ulo = (float)prm.pixel_border; c.uhi = (float)(prm.width - 1 - prm.pixel_border);
    c.vlo = (float)prm.pixel_border; c.vhi = (float)(prm.height - 1 - prm.pixel_border);
    c.z_eps = prm.z_eps;
  }

  float S[28], v[7];
#pragma unroll
  for (int l = 0; l < 28; ++l) S[l] = 0.f;
#pragma unroll
  for (int l = 0; l < 7; ++l) v[l] = 0.f;

  const float* __restrict__ Xi_base = Xs + (size_t)ix * N * 3;
  const float* __restrict__ Xj_base = Xs + (size_t)jx * N * 3;
  const float* __restrict__ Ci_base = Cs + (size_t)ix * N;
  const float* __restrict__ Cj_base = Cs + (size_t)jx * N;
  const int64_t* __restrict__ idx_e = idx + (size_t)e * N;
  const uint8_t* __restrict__ vm_e = valid_match + (size_t)e * N;
  const float* __restrict__ Q_e = Q + (size_t)e * N;

  const int k_begin = chunk * pts_per_cta;
  const int k_end = min(N, k_begin + pts_per_cta);
  for (int k0 = k_begin + threadIdx.x; k0 < k_end; k0 += kThreads * kUnroll) {
    // stage 1: streamed (read-once) per-edge data, evict-first
    long long ind[kUnroll];
    bool vm[kUnroll];
    float q[kUnroll], cj[kUnroll], Xj[kUnroll][3];
#pragma unroll
    for (int r = 0; r < kUnroll; ++r) {
      const int k = k0 + r * kThreads;
      const bool in = k < k_end;
      const int kk = in ? k : k_begin;
      vm[r] = in && (__ldcs(vm_e + kk) != 0);
      ind[r] = __ldcs(idx_e + kk);
      q[r] = __ldcs(Q_e + kk);
      cj[r] = __ldg(Cj_base + kk);
      Xj[r][0] = __ldg(Xj_base + (size_t)kk * 3 + 0);
      Xj[r][1] = __ldg(Xj_base + (size_t)kk * 3 + 1);
      Xj[r][2] = __ldg(Xj_base + (size_t)kk * 3 + 2);
      if (!vm[r]) ind[r] = 0;  // gn_kernels.cu:677
    }
    // stage 2: gathers from keyframe i
    float Xi[kUnroll][3], ci[kUnroll];
#pragma unroll
    for (int r = 0; r < kUnroll; ++r) {
      Xi[r][0] = __ldg(Xi_base + (size_t)ind[r] * 3 + 0);
      Xi[r][1] = __ldg(Xi_base + (size_t)ind[r] * 3 + 1);
      Xi[r][2] = __ldg(Xi_base + (size_t)ind[r] * 3 + 2);
      ci[r] = __ldg(Ci_base + ind[r]);
    }
#pragma unroll
    for (int r = 0; r < kUnroll; ++r)
      point_terms<MODE>(c, S, v, Xi[r], Xj[r], q[r], ci[r], cj[r], vm[r], ind[r]);
  }

  edge_epilogue<kThreads>(S, v, Ti, e, chunk, nchunks, nchunks, partials, tickets, edge_blocks,
                          peer);
}

// ---------------------------------------------------------------------------
// rays fast path kernel (v7): TWO points per thread in the two lanes of the packed fp32
// instructions (FFMA2 / FMUL2 / FADD2): lane x = point p, lane y = point p + 256 of a
// 512-point tile (so that the threads of a warp gather CONSECUTIVE points: two lanes then share
// a 32-byte sector and the request is coalesced before it reaches L2 -- with lanes on points
// 2k / 2k+1 every 16-byte gather was its own L2 sector request and the kernel ran into the L2
// request rate, profiles/r1_ncu_accumulate.md).  Every
// per-point quantity (transformed point, ray, residuals, Huber weights, Jacobian rows) and
// every accumulator is a float2, so the whole per-point arithmetic runs at half the
// instruction count of a scalar formulation and needs no pair-formation moves; the two lanes
// of an accumulator are added once, after the last tile.
// Inputs: the keyframe pointmaps repacked once per solve (frodo_gn_pack_points) in two layouts:
//   XC [K][N]      float4 (x,y,z,c)                    -- target of the data-dependent gather
//   XP [K][NB][256][8] (xA,xB,yA,yB,zA,zB,cA,cB)       -- streamed frame j, lanes pre-paired:
//      block b = 512 consecutive points, A = point p, B = point p + 256 of the block
//      (NB = ceil(N / 512), zero padded)
// Data movement per CTA (one edge, `pts_per_cta` points, tiles of 512 points = one pair per
// thread):
//   * the streamed arrays of a tile (idx 4 KB, XP of frame j 8 KB, Q 2 KB, valid 512 B:
//     contiguous, read once) are fetched by the TMA unit (cp.async.bulk, SASS UBLKCP) into a
//     kStages-deep shared-memory ring; a stage is handed back through an `empty` mbarrier on
//     which every thread arrives right after it has read its operands, so there is no
//     block-wide barrier in the loop;
//   * the frame-i gather of tile t + kGatherDepth is issued at the start of tile t as
//     asynchronous 16-byte copies global -> shared (cp.async, SASS LDGSTS), one commit group
//     per tile: no registers, no scoreboard (the in-register ring of v5 made every point wait
//     for a load issued a few instructions earlier, profiles/r1_ncu_accumulate.md).
// Requires N % 16 == 0, pts_per_cta % 512 == 0 and 16-byte aligned XC, idx, Q, valid.
// ---------------------------------------------------------------------------
// FRODO_ABL: development switches for ablation builds (profiles/exp_ablation.py); 0 in the product
#ifndef FRODO_ABL
#define FRODO_ABL 0
#endif
// L2 eviction priority of everything this kernel reads (cache-hint forms of the bulk copies and
// of the gathers; the policy is a kernel argument, see launch_gn_accumulate): a pass streams up to
// 2 GB through a 126 MB cache, and what it pushes out is the code and the tiles of the solve that
// follows, which then starts cold.  With evict-first on all reads, measured on the 64-keyframe
// graph: accumulate 473 -> 480 us (the gathers lose a little short-term reuse), Cholesky
// 138-141 -> 128-130 us; 32-keyframe calib graph 256 -> 260 and 83 -> 71 us.  Hinting only the
// streamed arrays, or those + XP, leaves the solve cold (the gathers alone flush the cache);
// gathers allocated in L1 as well: 518 us.
// (a warp-specialised persistent variant, measured slower -- 538 vs 473 us -- is kept as an
// experiment under profiles/experiments/gn_accumulate_ws.cuh, not in the product)
#ifndef FRODO_RAYS_THREADS
#define FRODO_RAYS_THREADS 128   // 4 CTAs x 128 threads per SM: 473 us; 2 x 256: 485 us; 8 x 64: 519 us
#endif
constexpr int kRaysThreads = FRODO_RAYS_THREADS;        // threads per CTA of the rays kernel
constexpr int kRaysCtasPerSM = 512 / kRaysThreads;      // 16 warps per SM either way
constexpr int kTilePts = 2 * kRaysThreads;  // one point pair per thread
constexpr int kStages = 5;
constexpr int kGatherDepth = 2;
constexpr int kOffIdx = 0;                       // int64  [512]
constexpr int kOffXP = kOffIdx + kTilePts * 8;   // float  [256][8] frame j, pair-interleaved
constexpr int kOffQ = kOffXP + kTilePts * 16;    // float  [512]
constexpr int kOffVm = kOffQ + kTilePts * 4;     // uint8  [512]
constexpr int kStageBytes = kOffVm + kTilePts;   // 29 bytes per point
static_assert(kStageBytes % 128 == 0, "stage alignment");
constexpr int kGatherBytes = kTilePts * 16;      // float4 [512] frame i (x,y,z,c)
constexpr int kRaysSmemBytes = kStages * kStageBytes + (kGatherDepth + 1) * kGatherBytes;

// XC (AoS float4) and XP (pair-interleaved per 512-point block) from Xs, Cs.  One thread per
// pair slot (keyframe f, block b, pair p): points n = 512 b + p and n + 256.
__global__ void pack_points_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs,
                                   float4* __restrict__ XC, float4* __restrict__ XP, int K, int N,
                                   int NB, const int32_t* __restrict__ need) {
  const size_t slots = (size_t)K * NB * kRaysThreads;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < slots; k += stride) {
    const int p = (int)(k % kRaysThreads);
    const size_t fb = k / kRaysThreads;
    const int b = (int)(fb % NB);
    const size_t f = fb / NB;
    const int want = need != nullptr ? __ldg(need + f) : 3;  // 1: XC row read, 2: XP row read
    if (want == 0) continue;
    const int nA = b * kTilePts + p, nB = nA + kRaysThreads;
    float4 A = make_float4(0.f, 0.f, 0.f, 0.f), Bv = A;
    if (nA < N) {
      const float* x = Xs + (f * N + nA) * 3;
      A = make_float4(__ldg(x), __ldg(x + 1), __ldg(x + 2), __ldg(Cs + f * N + nA));
      if (want & 1) XC[f * N + nA] = A;
    }
    if (nB < N) {
      const float* x = Xs + (f * N + nB) * 3;
      Bv = make_float4(__ldg(x), __ldg(x + 1), __ldg(x + 2), __ldg(Cs + f * N + nB));
      if (want & 1) XC[f * N + nB] = Bv;
    }
    if (want & 2) {
      XP[2 * k] = make_float4(A.x, Bv.x, A.y, Bv.y);
      XP[2 * k + 1] = make_float4(A.z, Bv.z, A.w, Bv.w);
    }
  }
}

// need[k] |= 1 if keyframe row k is the i of one of the E edges, |= 2 if it is the j.  One CTA.
__global__ void mark_keyframes_kernel(const int32_t* __restrict__ ii_row,
                                      const int32_t* __restrict__ jj_row, int E, int K,
                                      int32_t* __restrict__ need) {
  for (int k = threadIdx.x; k < K; k += blockDim.x) need[k] = 0;
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const int i = ii_row[e], j = jj_row[e];
    if (i >= 0 && i < K) atomicOr(need + i, 1);
    if (j >= 0 && j < K) atomicOr(need + j, 2);
  }
}

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }

// Huber weight of both lanes: min(1, 1.345 / (|e| sqrt(wq))) * wq ; wq = 0 -> 0.
__device__ __forceinline__ float2 huber_w2(float2 wq, float2 e) {
  const float2 t = mul2(mul2(wq, e), e);
  float2 h = mul2(bc(1.345f), f2(rsqrt_fast(t.x), rsqrt_fast(t.y)));
  h = f2(fminf(1.0f, h.x), fminf(1.0f, h.y));
  return mul2(h, wq);
}

// S += w x x^T, v += w x e over the columns in MASK (both lanes).
template <unsigned MASK>
__device__ __forceinline__ void pair_row(float2 (&S)[28], float2 (&v)[7], const float2 (&x)[7],
                                         float2 w, float2 e) {
#pragma unroll
  for (int m = 0; m < 7; ++m) {
    if (!((MASK >> m) & 1u)) continue;
    const float2 y = mul2(w, x[m]);
    v[m] = fma2(y, e, v[m]);
#pragma unroll
    for (int n = m; n < 7; ++n) {
      if (!((MASK >> n) & 1u)) continue;
      S[tri(m, n)] = fma2(y, x[n], S[tri(m, n)]);
    }
  }
}

// Two points (lanes x, y).  Same quantities as rays_point; the rotation columns carry their
// signs (c_0 = (0, r2, -r1), c_1 = (-r2, 0, r0), c_2 = (r1, -r0, 0)), so S and v are the
// canonical upper-triangular entries directly.  A ray row a of the local Jacobian is
//   x = [ d_a0 d_a1 d_a2 | c_a | 0 ],  d = (I - r r^T)/|P|,  c_a = row a of -[r]x,
// the distance row is [ r0 r1 r2 | 0 0 0 | |P| ]; 74 FFMA2 + 19 FMUL2 per pair of points.
__device__ __forceinline__ void rays_pair(const RaysConst& c, float2 (&S)[28], float2 (&v)[7],
                                          const float4 xiA, const float4 xiB, float2 xj0,
                                          float2 xj1, float2 xj2, float2 cj, float2 q, bool vmA,
                                          bool vmB) {
  const float2 P0 = fma2(bc(c.M[0]), xj0, fma2(bc(c.M[1]), xj1, fma2(bc(c.M[2]), xj2, bc(c.t[0]))));
  const float2 P1 = fma2(bc(c.M[3]), xj0, fma2(bc(c.M[4]), xj1, fma2(bc(c.M[5]), xj2, bc(c.t[1]))));
  const float2 P2 = fma2(bc(c.M[6]), xj0, fma2(bc(c.M[7]), xj1, fma2(bc(c.M[8]), xj2, bc(c.t[2]))));
  const float2 n2j = fma2(P0, P0, fma2(P1, P1, mul2(P2, P2)));
  const float2 a = f2(rsqrt_fast(n2j.x), rsqrt_fast(n2j.y));
  const float2 nj = mul2(n2j, a);
  const float2 r0 = mul2(P0, a), r1 = mul2(P1, a), r2 = mul2(P2, a);
  const float2 xi0 = f2(xiA.x, xiB.x), xi1 = f2(xiA.y, xiB.y), xi2 = f2(xiA.z, xiB.z);
  const float2 n2i = fma2(xi0, xi0, fma2(xi1, xi1, mul2(xi2, xi2)));
  const float2 nbi = f2(-rsqrt_fast(n2i.x), -rsqrt_fast(n2i.y));
  const float2 e0 = fma2(xi0, nbi, r0), e1 = fma2(xi1, nbi, r1), e2 = fma2(xi2, nbi, r2);
  const float2 e3 = fma2(n2i, nbi, nj);  // |P| - |Xi|
  const bool okA = vmA & (q.x > c.Q_thresh) & (xiA.w > c.C_thresh) & (cj.x > c.C_thresh);
  const bool okB = vmB & (q.y > c.Q_thresh) & (xiB.w > c.C_thresh) & (cj.y > c.C_thresh);
  const float2 wq = f2(okA ? q.x : 0.0f, okB ? q.y : 0.0f);
  const float2 wr = mul2(bc(c.sa2), wq), wd = mul2(bc(c.sb2), wq);
  const float2 w0 = huber_w2(wr, e0), w1 = huber_w2(wr, e1), w2 = huber_w2(wr, e2);
  const float2 w3 = huber_w2(wd, e3);
  // d r / d P = (I - r r^T) / |P| ; rotation columns -[r]x
  const float2 na = mul2(a, bc(-1.0f));
  const float2 s0 = mul2(na, r0), s1 = mul2(na, r1), s2 = mul2(na, r2);
  const float2 d01 = mul2(s0, r1), d02 = mul2(s0, r2), d12 = mul2(s1, r2);
  const float2 nr0 = mul2(r0, bc(-1.0f)), nr1 = mul2(r1, bc(-1.0f)), nr2 = mul2(r2, bc(-1.0f));
  const float2 z = f2(0.f, 0.f);
  {
    const float2 x[7] = {fma2(s0, r0, a), d01, d02, z, r2, nr1, z};
    pair_row<0x37u>(S, v, x, w0, e0);  // {0,1,2,4,5}
  }
  {
    const float2 x[7] = {d01, fma2(s1, r1, a), d12, nr2, z, r0, z};
    pair_row<0x2Fu>(S, v, x, w1, e1);  // {0,1,2,3,5}
  }
  {
    const float2 x[7] = {d02, d12, fma2(s2, r2, a), r1, nr0, z, z};
    pair_row<0x1Fu>(S, v, x, w2, e2);  // {0,1,2,3,4}
  }
  {
    const float2 x[7] = {r0, r1, r2, z, z, z, nj};
    pair_row<0x47u>(S, v, x, w3, e3);  // {0,1,2,6}
  }
}

// Points mode (gn_kernels.cu:2031-2158), two points per thread: e = P - Xi, three rows
// [e_a | -[P]x row a | P_a].
__device__ __forceinline__ void points_pair(const RaysConst& c, float2 (&S)[28], float2 (&v)[7],
                                            const float4 xiA, const float4 xiB, float2 xj0,
                                            float2 xj1, float2 xj2, float2 cj, float2 q, bool vmA,
                                            bool vmB) {
  const float2 P0 = fma2(bc(c.M[0]), xj0, fma2(bc(c.M[1]), xj1, fma2(bc(c.M[2]), xj2, bc(c.t[0]))));
  const float2 P1 = fma2(bc(c.M[3]), xj0, fma2(bc(c.M[4]), xj1, fma2(bc(c.M[5]), xj2, bc(c.t[1]))));
  const float2 P2 = fma2(bc(c.M[6]), xj0, fma2(bc(c.M[7]), xj1, fma2(bc(c.M[8]), xj2, bc(c.t[2]))));
  const float2 e0 = f2(P0.x - xiA.x, P0.y - xiB.x);
  const float2 e1 = f2(P1.x - xiA.y, P1.y - xiB.y);
  const float2 e2 = f2(P2.x - xiA.z, P2.y - xiB.z);
  const bool okA = vmA & (q.x > c.Q_thresh) & (xiA.w > c.C_thresh) & (cj.x > c.C_thresh);
  const bool okB = vmB & (q.y > c.Q_thresh) & (xiB.w > c.C_thresh) & (cj.y > c.C_thresh);
  const float2 ws = mul2(bc(c.sa2), f2(okA ? q.x : 0.0f, okB ? q.y : 0.0f));
  const float2 w0 = huber_w2(ws, e0), w1 = huber_w2(ws, e1), w2 = huber_w2(ws, e2);
  const float2 z = f2(0.f, 0.f), one = f2(1.f, 1.f);
  const float2 n0 = mul2(P0, bc(-1.0f)), n1 = mul2(P1, bc(-1.0f)), n2 = mul2(P2, bc(-1.0f));
  {
    const float2 x[7] = {one, z, z, z, P2, n1, P0};
    pair_row<0x71u>(S, v, x, w0, e0);  // {0,4,5,6}
  }
  {
    const float2 x[7] = {z, one, z, n2, z, P0, P1};
    pair_row<0x6Au>(S, v, x, w1, e1);  // {1,3,5,6}
  }
  {
    const float2 x[7] = {z, z, one, P1, n0, z, P2};
    pair_row<0x5Cu>(S, v, x, w2, e2);  // {2,3,4,6}
  }
}

// Calibrated mode (gn_kernels.cu:2410-2561), two points per thread: pixel residual of the
// projection of P against the target pixel (ind % width, ind / width), log-depth residual.
__device__ __forceinline__ void calib_pair(const RaysConst& c, float2 (&S)[28], float2 (&v)[7],
                                           const float4 xiA, const float4 xiB, float2 xj0,
                                           float2 xj1, float2 xj2, float2 cj, float2 q, bool vmA,
                                           bool vmB, unsigned indA, unsigned indB) {
  const float2 P0 = fma2(bc(c.M[0]), xj0, fma2(bc(c.M[1]), xj1, fma2(bc(c.M[2]), xj2, bc(c.t[0]))));
  const float2 P1 = fma2(bc(c.M[3]), xj0, fma2(bc(c.M[4]), xj1, fma2(bc(c.M[5]), xj2, bc(c.t[1]))));
  const float2 P2 = fma2(bc(c.M[6]), xj0, fma2(bc(c.M[7]), xj1, fma2(bc(c.M[8]), xj2, bc(c.t[2]))));
  const bool zA = (P2.x > c.z_eps) && (xiA.z > c.z_eps);
  const bool zB = (P2.y > c.z_eps) && (xiB.z > c.z_eps);
  // 1/z and log(z_j / z_i) through the special-function unit (1-2 ulp; the bar is 1e-4)
  const float2 zinv = f2(zA ? __fdividef(1.0f, P2.x) : 0.0f, zB ? __fdividef(1.0f, P2.y) : 0.0f);
  const float2 e2 = f2(zA ? __logf(__fdividef(P2.x, xiA.z)) : 0.0f,
                       zB ? __logf(__fdividef(P2.y, xiB.z)) : 0.0f);
  const float2 xz = mul2(P0, zinv), yz = mul2(P1, zinv);
  const float2 u = fma2(bc(c.fx), xz, bc(c.cx)), vv = fma2(bc(c.fy), yz, bc(c.cy));
  // ind = v_t * width + u_t: quotient estimate from floor(2^32 / width), one correction
  unsigned vtA = __umulhi(indA, c.width_rcp), vtB = __umulhi(indB, c.width_rcp);
  unsigned utA = indA - vtA * c.width, utB = indB - vtB * c.width;
  if (utA >= c.width) { utA -= c.width; ++vtA; }
  if (utB >= c.width) { utB -= c.width; ++vtB; }
  const float2 e0 = f2(u.x - (float)utA, u.y - (float)utB);
  const float2 e1 = f2(vv.x - (float)vtA, vv.y - (float)vtB);
  const bool inA = (u.x > c.ulo) && (u.x < c.uhi) && (vv.x > c.vlo) && (vv.x < c.vhi) && zA;
  const bool inB = (u.y > c.ulo) && (u.y < c.uhi) && (vv.y > c.vlo) && (vv.y < c.vhi) && zB;
  const bool okA = inA & vmA & (q.x > c.Q_thresh) & (xiA.w > c.C_thresh) & (cj.x > c.C_thresh);
  const bool okB = inB & vmB & (q.y > c.Q_thresh) & (xiB.w > c.C_thresh) & (cj.y > c.C_thresh);
  const float2 wq = f2(okA ? q.x : 0.0f, okB ? q.y : 0.0f);
  const float2 wp = mul2(bc(c.sa2), wq), wd = mul2(bc(c.sb2), wq);
  const float2 w0 = huber_w2(wp, e0), w1 = huber_w2(wp, e1), w2 = huber_w2(wd, e2);
  const float2 z = f2(0.f, 0.f), one = f2(1.f, 1.f);
  const float2 fxz = mul2(bc(c.fx), zinv), fyz = mul2(bc(c.fy), zinv);
  const float2 fxxz = mul2(bc(c.fx), xz), fyyz = mul2(bc(c.fy), yz);
  {
    // [fx/z, 0, -fx x/z^2, -fx xz yz, fx (1 + xz^2), -fx yz, 0]
    const float2 x[7] = {fxz, z, mul2(mul2(fxxz, zinv), bc(-1.0f)), mul2(mul2(fxxz, yz), bc(-1.0f)),
                         fma2(fxxz, xz, bc(c.fx)), mul2(mul2(bc(c.fx), yz), bc(-1.0f)), z};
    pair_row<0x3Du>(S, v, x, w0, e0);  // {0,2,3,4,5}
  }
  {
    // [0, fy/z, -fy y/z^2, -fy (1 + yz^2), fy xz yz, fy xz, 0]
    const float2 x[7] = {z, fyz, mul2(mul2(fyyz, zinv), bc(-1.0f)),
                         mul2(fma2(fyyz, yz, bc(c.fy)), bc(-1.0f)), mul2(fyyz, xz),
                         mul2(bc(c.fy), xz), z};
    pair_row<0x3Eu>(S, v, x, w1, e1);  // {1,2,3,4,5}
  }
  {
    const float2 x[7] = {z, z, zinv, yz, mul2(xz, bc(-1.0f)), z, one};
    pair_row<0x5Cu>(S, v, x, w2, e2);  // {2,3,4,6}
  }
}

#ifdef FRODO_ACC_STAMPS
__device__ unsigned long long g_acc_stamps[4 * 16384];
__device__ int g_acc_shift = 0;
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
__device__ __forceinline__ unsigned smid() {
  unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r;
}
#endif
template <int MODE>
__global__ void __launch_bounds__(kRaysThreads, kRaysCtasPerSM)
gn_accumulate_packed_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                            const float* __restrict__ Kmat,
                                 const float4* __restrict__ XC, const float4* __restrict__ XP,
                                 const int32_t* __restrict__ ii_row,
                                 const int32_t* __restrict__ jj_row,
                                 const int64_t* __restrict__ idx,
                                 const uint8_t* __restrict__ valid_match,
                                 const float* __restrict__ Q, int N, int pts_per_cta,
                                 float* __restrict__ partials, int32_t* __restrict__ tickets,
                                 double* __restrict__ edge_blocks,
                                 const int32_t* __restrict__ done, const GnPeerOut peer,
                                 int l2_evict_first) {
  if (done != nullptr && *done != 0) {
    peer_skip(peer);
    return;
  }
  const uint64_t l2_policy =
      l2_evict_first ? ptx::l2_policy_evict_first() : ptx::l2_policy_evict_normal();
  extern __shared__ __align__(128) unsigned char stage_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
#ifdef FRODO_ACC_STAMPS
  const int lin0 = blockIdx.y * gridDim.x + blockIdx.x;
  const int cta_lin = (lin0 + g_acc_shift) % (gridDim.x * gridDim.y);  // work item of this CTA
  const int e = cta_lin / gridDim.x;
  const int chunk = cta_lin % gridDim.x;
#else
  const int e = blockIdx.y;
  const int chunk = blockIdx.x;
#endif
  const int nchunks = gridDim.x;
  const int tid = threadIdx.x;
#ifdef FRODO_ACC_STAMPS
  if (tid == 32 && cta_lin < 16384) { g_acc_stamps[4 * cta_lin] = gtime(); g_acc_stamps[4 * cta_lin + 3] = smid(); }
#endif
  const int ix = ii_row[e], jx = jj_row[e];
  const int k_begin = chunk * pts_per_cta;
  const int k_end = min(N, k_begin + pts_per_cta);
  const int ntile = (k_end - k_begin + kTilePts - 1) / kTilePts;
  const float4* __restrict__ XCi = XC + (size_t)ix * N;
  const int NB = (N + kTilePts - 1) / kTilePts;
  const float4* gXPj = XP + (size_t)jx * NB * kTilePts;  // padded to whole blocks
  const int64_t* gIdx = idx + (size_t)e * N;
  const uint8_t* gVm = valid_match + (size_t)e * N;
  const float* gQ = Q + (size_t)e * N;

  // producer: one thread arms the stage barrier and issues the four bulk copies of tile t
  auto issue_tile = [&](int t) {
#if FRODO_ABL == 6
    const int k0 = k_begin + (t % kStages) * kTilePts;
#else
    const int k0 = k_begin + t * kTilePts;
#endif
    const uint32_t np = (uint32_t)min(kTilePts, k_end - k0);
    unsigned char* st = stage_raw + (size_t)(t % kStages) * kStageBytes;
    uint64_t* bar = &full_bar[t % kStages];
    ptx::mbar_arrive_expect_tx(bar, np * 13u + kTilePts * 16u);
    ptx::bulk_g2s_hint(st + kOffIdx, gIdx + k0, np * 8u, bar, l2_policy);
    ptx::bulk_g2s_hint(st + kOffXP, gXPj + k0, kTilePts * 16u, bar, l2_policy);  // whole (padded) block
    ptx::bulk_g2s_hint(st + kOffQ, gQ + k0, np * 4u, bar, l2_policy);
    ptx::bulk_g2s_hint(st + kOffVm, gVm + k0, np, bar, l2_policy);
  };
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], kRaysThreads);
    }
    ptx::mbar_fence_init();
#if FRODO_ABL != 4
    for (int t = 0; t < kStages && t < ntile; ++t) issue_tile(t);
#endif
  }

  const Sim3f Ti = load_pose(Twc, ix);
  const Sim3f Tj = load_pose(Twc, jx);
  RaysConst c;
  {
    const Sim3f Tij = rel_sim3(Ti, Tj);
    const float qx = Tij.q[0], qy = Tij.q[1], qz = Tij.q[2], qw = Tij.q[3], s = Tij.s;
    // same map as actSO3 (no renormalisation of q), times the relative scale
    c.M[0] = s * (1.f - 2.f * (qy * qy + qz * qz));
    c.M[1] = s * (2.f * (qx * qy - qz * qw));
    c.M[2] = s * (2.f * (qx * qz + qy * qw));
    c.M[3] = s * (2.f * (qx * qy + qz * qw));
    c.M[4] = s * (1.f - 2.f * (qx * qx + qz * qz));
    c.M[5] = s * (2.f * (qy * qz - qx * qw));
    c.M[6] = s * (2.f * (qx * qz - qy * qw));
    c.M[7] = s * (2.f * (qy * qz + qx * qw));
    c.M[8] = s * (1.f - 2.f * (qx * qx + qy * qy));
    c.t[0] = Tij.t[0]; c.t[1] = Tij.t[1]; c.t[2] = Tij.t[2];
    const float sa = 1.0f / prm.sigma_a;
    const float sb = (MODE == FRODO_GN_POINTS) ? 0.0f : 1.0f / prm.sigma_b;
    c.sa2 = sa * sa;
    c.sb2 = sb * sb;
    c.C_thresh = prm.C_thresh;
    c.Q_thresh = prm.Q_thresh;
    if (MODE == FRODO_GN_CALIB) {
      c.fx = __ldg(Kmat + 0); c.fy = __ldg(Kmat + 4); c.cx = __ldg(Kmat + 2); c.cy = __ldg(Kmat + 5);
      c.ulo = (float)prm.pixel_border; c.uhi = (float)(prm.width - 1 - prm.pixel_border);
      c.vlo = (float)prm.pixel_border; c.vhi = (float)(prm.height - 1 - prm.pixel_border);
      c.z_eps = prm.z_eps;
      c.width = prm.width > 1 ? (unsigned)prm.width : 1u;
      c.width_rcp = c.width > 1 ? (unsigned)(0x100000000ull / c.width) : 0xffffffffu;
    }
  }
  float2 S2[28], v2[7];
#pragma unroll
  for (int k = 0; k < 28; ++k) S2[k] = f2(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 7; ++k) v2[k] = f2(0.f, 0.f);
  __syncthreads();  // barriers initialised before anyone waits on them

  // Ring positions advance by increment (no modulo in the loop); everything below is
  // addressed through the shared window.
  const uint32_t sm_base = ptx::smem_u32(stage_raw);
  const uint32_t sm_gath = sm_base + kStages * kStageBytes + (uint32_t)tid * 32u;
  const uint32_t sm_full = ptx::smem_u32(&full_bar[0]);
  const uint32_t sm_empty = ptx::smem_u32(&empty_bar[0]);
  const int npts = k_end - k_begin;
  uint32_t full_phase = 0;  // bit s = parity to wait for on full_bar[s]
  // Tile with `left` points from its start on, in stage sg, gather buffer gg: wait until the
  // stage is full (every thread waits once per tile, here), then issue this thread's two
  // frame-i gathers as async copies.
  auto issue_gather = [&](int left, int sg, int gg) {
    ptx::mbar_wait(sm_full + 8u * sg, (full_phase >> sg) & 1u);
    full_phase ^= 1u << sg;
    if (tid < left) {
      const unsigned char* st = stage_raw + sg * kStageBytes;
      const long long* ip = reinterpret_cast<const long long*>(st + kOffIdx);
      const uint32_t dst = sm_gath + (uint32_t)gg * kGatherBytes;
      // An unmatched point has weight zero whatever frame-i point stands beside it: the reference
      // reads point 0 there (gn_kernels.cu:677); here nothing is loaded and the slot keeps what
      // it held (the constant written below, or an earlier matched point: finite either way).
      // Every CTA of keyframe i fetching the same 16 bytes is a hot spot in one L2 slice.
      if (st[kOffVm + tid] != 0) {
#if FRODO_ABL == 5
        ptx::cp_async_16(dst, XCi + (k_end - left + tid));
#else
        ptx::cp_async_16_hint(dst, XCi + ip[tid], l2_policy);
#endif
      }
      if (tid + kRaysThreads < left && st[kOffVm + tid + kRaysThreads] != 0) {
#if FRODO_ABL == 5
        ptx::cp_async_16(dst + 16u, XCi + (k_end - left + tid + kRaysThreads));
#else
        ptx::cp_async_16_hint(dst + 16u, XCi + ip[tid + kRaysThreads], l2_policy);
#endif
      }
    }
  };

  {  // this thread's slots of every gather buffer: a finite stand-in for unmatched points
    float4* slot = reinterpret_cast<float4*>(stage_raw + kStages * kStageBytes) + 2 * tid;
#pragma unroll
    for (int g = 0; g <= kGatherDepth; ++g) {
      slot[g * (kGatherBytes / 16)] = make_float4(1.f, 1.f, 1.f, 0.f);
      slot[g * (kGatherBytes / 16) + 1] = make_float4(1.f, 1.f, 1.f, 0.f);
    }
  }
#if FRODO_ABL != 4
#pragma unroll
  for (int g = 0; g < kGatherDepth; ++g) {
    if (g < ntile) issue_gather(npts - g * kTilePts, g % kStages, g);
    ptx::cp_async_commit();
  }
#endif
  uint32_t empty_phase = 0;          // producer only
  int s_cur = 0, g_cur = 0;          // stage / gather buffer of tile t
  int s_ahead = kGatherDepth % kStages, g_ahead = kGatherDepth;  // of tile t + kGatherDepth
  int s_prev = kStages - 1;          // stage of tile t - 1
  int left = npts;                   // points from the start of tile t to the end of the span
  for (int t = 0; t < ntile; ++t) {
#if FRODO_ABL != 4
    if (tid == 0 && t >= 1 && t - 1 + kStages < ntile) {
      // every thread has read its operands of tile t-1: refill that stage
      ptx::mbar_wait(sm_empty + 8u * s_prev, (empty_phase >> s_prev) & 1u);
      empty_phase ^= 1u << s_prev;
      issue_tile(t - 1 + kStages);
    }
    if (t + kGatherDepth < ntile) issue_gather(left - kGatherDepth * kTilePts, s_ahead, g_ahead);
    ptx::cp_async_commit();
    ptx::cp_async_wait<kGatherDepth>();  // this thread's gathers of tile t have landed
#endif
    if (tid < left) {
      const unsigned char* st = stage_raw + s_cur * kStageBytes;
      float4 ja = reinterpret_cast<const float4*>(st + kOffXP)[2 * tid];
      float4 jb = reinterpret_cast<const float4*>(st + kOffXP)[2 * tid + 1];
      const float* qp = reinterpret_cast<const float*>(st + kOffQ);
      const float4* gb = reinterpret_cast<const float4*>(
          stage_raw + kStages * kStageBytes + g_cur * kGatherBytes) + 2 * tid;
      const float4 xiA = gb[0];
      float4 xiB;
      float2 q;
      bool vmA = st[kOffVm + tid] != 0, vmB;
      unsigned indA = 0, indB = 0;  // target pixel of the match (calib), low word of idx
      if (MODE == FRODO_GN_CALIB)
        indA = reinterpret_cast<const unsigned*>(st + kOffIdx)[2 * tid];
      if (tid + kRaysThreads < left) {
        xiB = gb[1];
        q = f2(qp[tid], qp[tid + kRaysThreads]);
        vmB = st[kOffVm + tid + kRaysThreads] != 0;
        if (MODE == FRODO_GN_CALIB)
          indB = reinterpret_cast<const unsigned*>(st + kOffIdx)[2 * (tid + kRaysThreads)];
      } else {  // ragged tail: lane y repeats lane x with zero weight
        xiB = xiA;
        q = f2(qp[tid], qp[tid]);
        vmB = false;
        ja = make_float4(ja.x, ja.x, ja.z, ja.z);
        jb = make_float4(jb.x, jb.x, jb.z, jb.z);
      }
      ptx::mbar_arrive(sm_empty + 8u * s_cur);  // operands are in registers
      if (MODE == FRODO_GN_RAYS)
        rays_pair(c, S2, v2, xiA, xiB, f2(ja.x, ja.y), f2(ja.z, ja.w), f2(jb.x, jb.y),
                  f2(jb.z, jb.w), q, vmA, vmB);
      else if (MODE == FRODO_GN_POINTS)
        points_pair(c, S2, v2, xiA, xiB, f2(ja.x, ja.y), f2(ja.z, ja.w), f2(jb.x, jb.y),
                    f2(jb.z, jb.w), q, vmA, vmB);
      else
        calib_pair(c, S2, v2, xiA, xiB, f2(ja.x, ja.y), f2(ja.z, ja.w), f2(jb.x, jb.y),
                   f2(jb.z, jb.w), q, vmA, vmB, vmA ? indA : 0u, vmB ? indB : 0u);
    } else {
      ptx::mbar_arrive(sm_empty + 8u * s_cur);
    }
    s_prev = s_cur;
    s_cur = s_cur + 1 == kStages ? 0 : s_cur + 1;
    s_ahead = s_ahead + 1 == kStages ? 0 : s_ahead + 1;
    g_cur = g_cur == kGatherDepth ? 0 : g_cur + 1;
    g_ahead = g_ahead == kGatherDepth ? 0 : g_ahead + 1;
    left -= kTilePts;
  }
  float S[28], v[7];
#pragma unroll
  for (int k = 0; k < 28; ++k) S[k] = S2[k].x + S2[k].y;
#pragma unroll
  for (int k = 0; k < 7; ++k) v[k] = v2[k].x + v2[k].y;
#ifdef FRODO_ACC_STAMPS
  if (tid == 32 && cta_lin < 16384) g_acc_stamps[4 * cta_lin + 1] = gtime();
#endif
  edge_epilogue<kRaysThreads>(S, v, Ti, e, chunk, nchunks, nchunks, partials, tickets,
                              edge_blocks, peer);
#ifdef FRODO_ACC_STAMPS
  if (tid == 32 && cta_lin < 16384) g_acc_stamps[4 * cta_lin + 2] = gtime();
#endif
}


}  // namespace

// Points per accumulate CTA, from a model of how the grid fills the device: `full` rounds in
// which every SM holds kRaysCtasPerSM CTAs, then a last round with r = 1..4 CTAs per SM, where a
// CTA sharing its SM with fewer neighbours runs faster (g(r), measured: a 3-CTA SM finishes in
// 0.8 of the time of a 4-CTA one, 2 CTAs in 0.645, 1 in 0.53).  Each CTA costs its points plus
// ~400 points of pipeline fill and reduction epilogue.  Fitted to the sweep in
// profiles/r2_points_per_cta.txt (E = 24 ... 256 at 512x384); picks within 2 % of the best there.
int gn_pts_per_cta(int E, int N) {
  static const int forced = [] {  // development override
    const char* e = getenv("FRODO_GN_PPC");
    return e ? atoi(e) : 0;
  }();
  if (forced > 0) return forced;
  static_assert(kRaysCtasPerSM == 4, "g() below is measured for four CTAs per SM");
  const double g[5] = {0.0, 0.53, 0.645, 0.8, 1.0};
  const long long cand[] = {2048, 3072, 4096, 6144, 8192, 12288, 16384, 24576, 32768};
  const long long sms = num_sms(), slots = (long long)kRaysCtasPerSM * sms;
  long long best = 2048;
  double best_cost = 1e300;
  for (long long ppc : cand) {
    const long long nch = ((long long)(N > 0 ? N : 1) + ppc - 1) / ppc;
    const long long ncta = nch * (long long)(E > 0 ? E : 1);
    const long long full = ncta / slots, rem = ncta - full * slots;
    const long long r = (rem + sms - 1) / sms;  // CTAs per SM in the last round
    const long long eff = ppc < N ? ppc : ((N + kTilePts - 1) / kTilePts) * kTilePts;
    const double cost = ((double)eff + 400.0) * ((double)full + g[r]);
    if (cost < best_cost) {
      best_cost = cost;
      best = ppc;
    }
  }
  return (int)best;
}

#ifdef FRODO_ACC_STAMPS
extern "C" int frodo_debug_acc_shift(int shift) {
  return (int)cudaMemcpyToSymbol(g_acc_shift, &shift, sizeof(int));
}
extern "C" int frodo_debug_acc_stamps(unsigned long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_acc_stamps, (size_t)n * 4 * 8);
}
#endif
size_t gn_packed_bytes(int K, int N) {
  const size_t k = (size_t)(K > 0 ? K : 1), n = (size_t)(N > 0 ? N : 1);
  const size_t nb = (n + kTilePts - 1) / kTilePts;
  return k * n * 16 + k * nb * kTilePts * 16;  // XC, then XP padded to whole blocks
}

int launch_gn_mark_keyframes(const int32_t* ii_row, const int32_t* jj_row, int E, int K,
                             int32_t* need, cudaStream_t s) {
  if (K <= 0) return FRODO_OK;
  mark_keyframes_kernel<<<1, 256, 0, s>>>(ii_row, jj_row, E, K, need);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_gn_pack_points(const float* Xs, const float* Cs, int K, int N, float* XC,
                          cudaStream_t s, const int32_t* need) {
  const size_t total = (size_t)K * (size_t)N;
  if (total == 0) return FRODO_OK;
  const int NB = (N + kTilePts - 1) / kTilePts;
  const size_t slots = (size_t)K * NB * kRaysThreads;
  const size_t want = (slots + 255) / 256;
  const size_t resident = (size_t)8 * (size_t)num_sms();
  const int blocks = (int)(want < resident ? want : resident);
  float4* xc = reinterpret_cast<float4*>(XC);
  pack_points_kernel<<<blocks, 256, 0, s>>>(Xs, Cs, xc, xc + total, K, N, NB, need);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_gn_pack_slot(const float* X, const float* C, int K_total, int slot, int N, float* XC,
                        cudaStream_t s) {
  if (N == 0) return FRODO_OK;
  const int NB = (N + kTilePts - 1) / kTilePts;
  const size_t slots = (size_t)NB * kRaysThreads;
  const int blocks = (int)((slots + 255) / 256);
  float4* xc = reinterpret_cast<float4*>(XC);
  float4* xc_row = xc + (size_t)slot * N;
  float4* xp_row = xc + (size_t)K_total * N + (size_t)slot * NB * kRaysThreads * 2;
  pack_points_kernel<<<blocks, 256, 0, s>>>(X, C, xc_row, xp_row, 1, N, NB, nullptr);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_gn_accumulate(const frodo_gn_params& prm, const float* Twc, const float* Xs,
                         const float* Cs, const float* Kmat, const float* XC,
                         const int32_t* ii_row, const int32_t* jj_row, const int64_t* idx,
                         const uint8_t* valid, const float* Q, int K, int N, int E,
                         float* partials, int32_t* tickets, double* edge_blocks,
                         const int32_t* done, cudaStream_t s, const GnPeerOut* peer_out) {
  if (E == 0) return FRODO_OK;
  GnPeerOut peer = {};
  if (peer_out != nullptr) peer = *peer_out;
  const int ppc = gn_pts_per_cta(E, N);
  const int nchunks = (N + ppc - 1) / ppc;
  dim3 grid(nchunks > 0 ? nchunks : 1, E);
  const bool aligned16 =
      ((reinterpret_cast<uintptr_t>(XC) | reinterpret_cast<uintptr_t>(idx) |
        reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(valid)) & 15) == 0;
  // packed fast path (all three modes): needs the repacked pointmaps and 16-byte alignment;
  // calib additionally needs N < 2^32 target pixels, which N as an int guarantees
  if (XC != nullptr && (N % 16) == 0 && aligned16 &&
      (prm.mode != FRODO_GN_CALIB || Kmat != nullptr)) {
    static PerDeviceOnce once;
    int dev;
    const size_t smem = (size_t)kRaysSmemBytes;
    if (once.needed(&dev)) {
      FRODO_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_packed_kernel<FRODO_GN_RAYS>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      FRODO_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_packed_kernel<FRODO_GN_POINTS>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      FRODO_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_packed_kernel<FRODO_GN_CALIB>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      once.mark(dev);
    }
    const float4* xc = reinterpret_cast<const float4*>(XC);
    const float4* xp = xc + (size_t)K * N;
    // Evict-first reads keep the following solve warm at a ~1.4 % cost to this kernel: worth it
    // when the solve is a large share of the iteration (smaller graphs, slices of a sharded
    // solve), about even on the 64-keyframe graph on one GPU, where it is left off.
    const int ef = ((long long)E * (long long)N <= (32LL << 20)) ? 1 : 0;
    if (prm.mode == FRODO_GN_RAYS)
      gn_accumulate_packed_kernel<FRODO_GN_RAYS><<<grid, kRaysThreads, smem, s>>>(
          prm, Twc, Kmat, xc, xp, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer, ef);
    else if (prm.mode == FRODO_GN_POINTS)
      gn_accumulate_packed_kernel<FRODO_GN_POINTS><<<grid, kRaysThreads, smem, s>>>(
          prm, Twc, Kmat, xc, xp, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer, ef);
    else
      gn_accumulate_packed_kernel<FRODO_GN_CALIB><<<grid, kRaysThreads, smem, s>>>(
          prm, Twc, Kmat, xc, xp, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer, ef);
    FRODO_CUDA_TRY(cudaGetLastError());
    return FRODO_OK;
  }
  switch (prm.mode) {
    case FRODO_GN_RAYS:
      gn_accumulate_kernel<FRODO_GN_RAYS><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer);
      break;
    case FRODO_GN_POINTS:
      gn_accumulate_kernel<FRODO_GN_POINTS><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer);
      break;
    case FRODO_GN_CALIB:
      gn_accumulate_kernel<FRODO_GN_CALIB><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done, peer);
      break;
    default:
      set_error("unknown GN mode %d", prm.mode);
      return FRODO_ERR_ARG;
  }
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
