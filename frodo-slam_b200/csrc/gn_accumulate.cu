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
// rays, fast path.  The weighted outer products run on the packed fp32 pipe (FFMA2, two
// fp32 lanes per issue slot).  A ray row a (a = 0,1,2) of the local Jacobian is
//   x = [ d_a0 d_a1 d_a2 | c_a ],  d = (I - r r^T)/|P|,  c_a = row a of -[r]x
// and c_a has exactly two non-zeros (+-r_p, +-r_q).  With the residual e_a appended it is
// packed into three register pairs without padding:
//   T0 = (d_a0, d_a1)   T1 = (d_a2, e_a)   M = (r_p, r_q)   (signs applied at unpack)
// and adds, with W* = w * pair:
//   TT  = W0.x*T0, W0.y*T0          -> S00 S01 (S10) S11        shared by all rows
//   TU  = W0.x*T1, W0.y*T1          -> S02 v0 S12 v1             shared
//   UU  = W1.x*T1                   -> S22 v2                    shared
//   TM  = W0.x*M, W0.y*M            -> S0p S0q S1p S1q           per row
//   UM  = W1.x*M, W1.y*M            -> S2p S2q vp vq             per row
//   MM  = WM.x*M, WM.y*M            -> Spp Spq (Sqp) Sqq         per row
// i.e. 11 FFMA2 + 3 FMUL2 per row.  The distance row [r0 r1 r2 | nj | e3] uses the same
// T0/T1 accumulators with T0 = (r0,r1), T1 = (r2,e3) plus the nj column.
// Weights: w = huber(|e| sqrt(q)/sigma) q/sigma^2 = min(1, 1.345 rsqrt(t)) wq with
// t = wq e^2, wq = q/sigma^2 (same value as gn_kernels.cu:214-217,730-741, no sqrt(q)).
// ---------------------------------------------------------------------------
struct RaysAcc {
  float2 TTa, TTb, TUa, TUb, UU;  // shared T-block accumulators
  float2 TMa[3], TMb[3], UMa[3], UMb[3], MMa[3], MMb[3];  // per ray row
  float2 N6a, N6b;  // (S06, S16), (S26, v6)
  float n66;
};

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 bc(float a) { return make_float2(a, a); }

__device__ __forceinline__ float rsqrt_fast(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float huber_w(float wq, float e) {
  // min(1, 1.345 / (|e| sqrt(wq))) * wq ; wq = 0 -> 0
  const float t = wq * e * e;
  return fminf(1.0f, 1.345f * rsqrt_fast(t)) * wq;
}

__device__ __forceinline__ void t_block(RaysAcc& A, float2 W0, float2 W1, float2 T0, float2 T1) {
  A.TTa = __ffma2_rn(bc(W0.x), T0, A.TTa);
  A.TTb = __ffma2_rn(bc(W0.y), T0, A.TTb);
  A.TUa = __ffma2_rn(bc(W0.x), T1, A.TUa);
  A.TUb = __ffma2_rn(bc(W0.y), T1, A.TUb);
  A.UU = __ffma2_rn(bc(W1.x), T1, A.UU);
}

template <int ROW>
__device__ __forceinline__ void rays_row(RaysAcc& A, float w, float2 T0, float2 T1, float2 M) {
  const float2 W0 = __fmul2_rn(bc(w), T0);
  const float2 W1 = __fmul2_rn(bc(w), T1);
  const float2 WM = __fmul2_rn(bc(w), M);
  t_block(A, W0, W1, T0, T1);
  A.TMa[ROW] = __ffma2_rn(bc(W0.x), M, A.TMa[ROW]);
  A.TMb[ROW] = __ffma2_rn(bc(W0.y), M, A.TMb[ROW]);
  A.UMa[ROW] = __ffma2_rn(bc(W1.x), M, A.UMa[ROW]);
  A.UMb[ROW] = __ffma2_rn(bc(W1.y), M, A.UMb[ROW]);
  A.MMa[ROW] = __ffma2_rn(bc(WM.x), M, A.MMa[ROW]);
  A.MMb[ROW] = __ffma2_rn(bc(WM.y), M, A.MMb[ROW]);
}

struct RaysConst {
  float M[9];  // s_ij R(q_ij)
  float t[3];
  float sa2, sb2, C_thresh, Q_thresh;
};

__device__ __forceinline__ void rays_point(const RaysConst& c, RaysAcc& A, float xi0, float xi1,
                                           float xi2, float xj0, float xj1, float xj2, float q,
                                           float ci, float cj, bool vm) {
  const float P0 = fmaf(c.M[0], xj0, fmaf(c.M[1], xj1, fmaf(c.M[2], xj2, c.t[0])));
  const float P1 = fmaf(c.M[3], xj0, fmaf(c.M[4], xj1, fmaf(c.M[5], xj2, c.t[1])));
  const float P2 = fmaf(c.M[6], xj0, fmaf(c.M[7], xj1, fmaf(c.M[8], xj2, c.t[2])));
  const float n2j = fmaf(P0, P0, fmaf(P1, P1, P2 * P2));
  const float a = rsqrt_fast(n2j);
  const float nj = n2j * a;
  const float r0 = P0 * a, r1 = P1 * a, r2 = P2 * a;
  const float n2i = fmaf(xi0, xi0, fmaf(xi1, xi1, xi2 * xi2));
  const float bi = rsqrt_fast(n2i);
  const float ni = n2i * bi;
  const float e0 = fmaf(-xi0, bi, r0);
  const float e1 = fmaf(-xi1, bi, r1);
  const float e2 = fmaf(-xi2, bi, r2);
  const float e3 = nj - ni;
  const bool valid = vm & (q > c.Q_thresh) & (ci > c.C_thresh) & (cj > c.C_thresh);
  const float wq = valid ? q : 0.0f;
  const float wr = c.sa2 * wq, wd = c.sb2 * wq;
  const float w0 = huber_w(wr, e0), w1 = huber_w(wr, e1), w2 = huber_w(wr, e2);
  const float w3 = huber_w(wd, e3);
  // d r / d P = (I - r r^T) / |P|
  const float s0 = -a * r0, s1 = -a * r1, s2 = -a * r2;
  const float d00 = fmaf(s0, r0, a), d11 = fmaf(s1, r1, a), d22 = fmaf(s2, r2, a);
  const float d01 = s0 * r1, d02 = s0 * r2, d12 = s1 * r2;
  // c_0 = (0, r2, -r1)  c_1 = (-r2, 0, r0)  c_2 = (r1, -r0, 0): magnitudes only, signs at unpack
  rays_row<0>(A, w0, f2(d00, d01), f2(d02, e0), f2(r1, r2));
  rays_row<1>(A, w1, f2(d01, d11), f2(d12, e1), f2(r0, r2));
  rays_row<2>(A, w2, f2(d02, d12), f2(d22, e2), f2(r0, r1));
  // distance row [r0 r1 r2 | nj | e3]
  {
    const float2 T0 = f2(r0, r1), T1 = f2(r2, e3);
    const float2 W0 = __fmul2_rn(bc(w3), T0), W1 = __fmul2_rn(bc(w3), T1);
    t_block(A, W0, W1, T0, T1);
    A.N6a = __ffma2_rn(W0, bc(nj), A.N6a);
    A.N6b = __ffma2_rn(W1, bc(nj), A.N6b);
    A.n66 = fmaf(w3 * nj, nj, A.n66);
  }
}

__host__ __device__ constexpr int tri(int n, int m) { return n * 7 - n * (n - 1) / 2 + (m - n); }

// Folds the packed accumulators into the canonical 28 upper-triangular entries + 7.
// Row a's rotation entries sit at columns 3..5 with
//   c_0: col 4 = +r2, col 5 = -r1 ; c_1: col 3 = -r2, col 5 = +r0 ; c_2: col 3 = +r1, col 4 = -r0
// and the packed pairs hold M_0 = (r1, r2), M_1 = (r0, r2), M_2 = (r0, r1).
__device__ __forceinline__ void rays_unpack(const RaysAcc& A, float (&S)[28], float (&v)[7]) {
  S[tri(0, 0)] = A.TTa.x;
  S[tri(0, 1)] = A.TTa.y;
  S[tri(1, 1)] = A.TTb.y;
  S[tri(0, 2)] = A.TUa.x;
  S[tri(1, 2)] = A.TUb.x;
  S[tri(2, 2)] = A.UU.x;
  v[0] = A.TUa.y;
  v[1] = A.TUb.y;
  v[2] = A.UU.y;
  // T x M: S[n][col]
  //   row0: r1 -> col5 (-), r2 -> col4 (+) ; row1: r0 -> col5 (+), r2 -> col3 (-)
  //   row2: r0 -> col4 (-), r1 -> col3 (+)
  S[tri(0, 3)] = -A.TMa[1].y + A.TMa[2].y;
  S[tri(0, 4)] = A.TMa[0].y - A.TMa[2].x;
  S[tri(0, 5)] = -A.TMa[0].x + A.TMa[1].x;
  S[tri(1, 3)] = -A.TMb[1].y + A.TMb[2].y;
  S[tri(1, 4)] = A.TMb[0].y - A.TMb[2].x;
  S[tri(1, 5)] = -A.TMb[0].x + A.TMb[1].x;
  S[tri(2, 3)] = -A.UMa[1].y + A.UMa[2].y;
  S[tri(2, 4)] = A.UMa[0].y - A.UMa[2].x;
  S[tri(2, 5)] = -A.UMa[0].x + A.UMa[1].x;
  v[3] = -A.UMb[1].y + A.UMb[2].y;
  v[4] = A.UMb[0].y - A.UMb[2].x;
  v[5] = -A.UMb[0].x + A.UMb[1].x;
  // M x M
  S[tri(3, 3)] = A.MMb[1].y + A.MMb[2].y;   // r2 r2 (row1) + r1 r1 (row2)
  S[tri(4, 4)] = A.MMb[0].y + A.MMa[2].x;   // r2 r2 (row0) + r0 r0 (row2)
  S[tri(5, 5)] = A.MMa[0].x + A.MMa[1].x;   // r1 r1 (row0) + r0 r0 (row1)
  S[tri(3, 4)] = -A.MMa[2].y;               // row2: (+r1)(-r0)
  S[tri(3, 5)] = -A.MMa[1].y;               // row1: (+r0)(-r2)
  S[tri(4, 5)] = -A.MMa[0].y;               // row0: (-r1)(+r2)
  S[tri(0, 6)] = A.N6a.x;
  S[tri(1, 6)] = A.N6a.y;
  S[tri(2, 6)] = A.N6b.x;
  S[tri(3, 6)] = 0.f;
  S[tri(4, 6)] = 0.f;
  S[tri(5, 6)] = 0.f;
  S[tri(6, 6)] = A.n66;
  v[6] = A.N6b.y;
}

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
template <int THREADS>
__device__ __forceinline__ void edge_epilogue(float (&S)[28], float (&v)[7], const Sim3f& Ti,
                                              int e, int chunk, int nchunks, int stride,
                                              float* __restrict__ partials,
                                              int32_t* __restrict__ tickets,
                                              double* __restrict__ edge_blocks) {
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
  __syncthreads();
  float* __restrict__ my_part = partials + ((size_t)e * stride + chunk) * kGnBlock;
  if (threadIdx.x < kGnBlock) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) x += warp_part[w][threadIdx.x];
    my_part[threadIdx.x] = x;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int t = atomicAdd(tickets + e, 1);
    is_last = (t == nchunks - 1);
  }
  __syncthreads();
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
  if (threadIdx.x == 64) adjoint_matrix(Ti, Md);
  if (threadIdx.x == 0) tickets[e] = 0;  // ready for the next iteration
  __syncthreads();
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    const int a = min(r, cc), b = max(r, cc);
    Sfull[threadIdx.x] = Sd[a * 7 - a * (a - 1) / 2 + (b - a)];
  }
  __syncthreads();
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    double acc = 0.0;
    for (int a = 0; a < 7; ++a) acc += Md[r * 7 + a] * Sfull[a * 7 + cc];
    MS[threadIdx.x] = acc;
  }
  __syncthreads();
  double* __restrict__ out = edge_blocks + (size_t)e * kGnBlock;
  if (threadIdx.x < 49) {
    const int r = threadIdx.x / 7, cc = threadIdx.x % 7;
    if (cc >= r) {
      double acc = 0.0;
      for (int b = 0; b < 7; ++b) acc += MS[r * 7 + b] * Md[cc * 7 + b];
      out[r * 7 - r * (r - 1) / 2 + (cc - r)] = acc;
    }
  } else if (threadIdx.x < 56) {
    const int r = threadIdx.x - 49;
    double acc = 0.0;
    for (int a = 0; a < 7; ++a) acc += Md[r * 7 + a] * Sd[28 + a];
    out[28 + r] = acc;
  }
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 2)
gn_accumulate_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                     const float* __restrict__ Xs, const float* __restrict__ Cs,
                     const float* __restrict__ Kmat, const int32_t* __restrict__ ii_row,
                     const int32_t* __restrict__ jj_row, const int64_t* __restrict__ idx,
                     const uint8_t* __restrict__ valid_match, const float* __restrict__ Q, int N,
                     int pts_per_cta, float* __restrict__ partials, int32_t* __restrict__ tickets,
                     double* __restrict__ edge_blocks, const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
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

  edge_epilogue<kThreads>(S, v, Ti, e, chunk, nchunks, nchunks, partials, tickets, edge_blocks);
}

// ---------------------------------------------------------------------------
// rays fast path kernel: see the comment above RaysAcc for the arithmetic.
// Inputs: the keyframe pointmaps repacked once per solve into XC [K][N] float4 (x,y,z,c)
// (frodo_gn_pack_points), so that the data-dependent gather of frame i is ONE aligned
// 16-byte load per point instead of four 4-byte loads (4x fewer L1 requests and sectors:
// the unpacked gather ran at ~23 sectors per warp request and kept the L1 tag stage busy
// for the whole kernel, profiles/r1_ncu_accumulate.md).
// Data movement per CTA (one edge, `pts_per_cta` points, tiles of 768 points):
//   * the streamed arrays of a tile (idx 6 KB, XC of frame j 12 KB, Q 3 KB, valid 768 B:
//     contiguous, read once) are fetched by the TMA unit (cp.async.bulk, SASS UBLKCP) into a
//     3-stage shared-memory ring, one transaction mbarrier per stage;
//   * lanes work on consecutive points (thread t: points t, t+256, t+512 of the tile), so a
//     gather request covers 16 consecutive-ish sectors;
//   * the gather of tile t+1 is issued at the start of tile t as asynchronous 16-byte copies
//     global -> shared (cp.async, SASS LDGSTS) into a double-buffered gather area, one commit
//     group per tile.  It needs no registers and no scoreboard: with the earlier in-register
//     ring every refill shared a scoreboard with the loads still in flight, so each point
//     waited for a gather issued a few instructions before (35 % of all stall samples,
//     profiles/r1_ncu_accumulate.md v5 -> v6);
//   * a stage is refilled with tile t+3 after the block-wide barrier that follows its last
//     reader.
// Requires N % 16 == 0, pts_per_cta % 768 == 0 and 16-byte aligned XC, idx, Q, valid.
// ---------------------------------------------------------------------------
constexpr int kPtsPerThread = 3;
constexpr int kTilePts = kPtsPerThread * kThreads;  // 768
constexpr int kStages = 3;
constexpr int kOffIdx = 0;                       // int64  [768]
constexpr int kOffXC = kOffIdx + kTilePts * 8;   // float4 [768]   frame j (x,y,z,c)
constexpr int kOffQ = kOffXC + kTilePts * 16;    // float  [768]
constexpr int kOffVm = kOffQ + kTilePts * 4;     // uint8  [768]
constexpr int kStageBytes = kOffVm + kTilePts;   // 29 bytes per point
static_assert(kStageBytes % 128 == 0, "stage alignment");
constexpr int kGatherBytes = kTilePts * 16;    // float4 [768] frame i (x,y,z,c), x2 buffers
constexpr int kRaysSmemBytes = kStages * kStageBytes + 2 * kGatherBytes;

__global__ void pack_points_kernel(const float* __restrict__ Xs, const float* __restrict__ Cs,
                                   float4* __restrict__ XC, size_t total) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += stride) {
    const float* x = Xs + p * 3;
    XC[p] = make_float4(__ldg(x), __ldg(x + 1), __ldg(x + 2), __ldg(Cs + p));
  }
}

__global__ void __launch_bounds__(kThreads, 2)
gn_accumulate_rays_packed_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                                 const float4* __restrict__ XC,
                                 const int32_t* __restrict__ ii_row,
                                 const int32_t* __restrict__ jj_row,
                                 const int64_t* __restrict__ idx,
                                 const uint8_t* __restrict__ valid_match,
                                 const float* __restrict__ Q, int N, int pts_per_cta,
                                 float* __restrict__ partials, int32_t* __restrict__ tickets,
                                 double* __restrict__ edge_blocks,
                                 const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  extern __shared__ __align__(128) unsigned char stage_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  const int e = blockIdx.y;
  const int chunk = blockIdx.x;
  const int nchunks = gridDim.x;
  const int tid = threadIdx.x;
  const int ix = ii_row[e], jx = jj_row[e];
  const int k_begin = chunk * pts_per_cta;
  const int k_end = min(N, k_begin + pts_per_cta);
  const int ntile = (k_end - k_begin + kTilePts - 1) / kTilePts;
  const float4* __restrict__ XCi = XC + (size_t)ix * N;
  const float4* gXCj = XC + (size_t)jx * N;
  const int64_t* gIdx = idx + (size_t)e * N;
  const uint8_t* gVm = valid_match + (size_t)e * N;
  const float* gQ = Q + (size_t)e * N;

  // producer: one thread arms the stage barrier and issues the four bulk copies of tile t
  auto issue_tile = [&](int t) {
    const int k0 = k_begin + t * kTilePts;
    const uint32_t np = (uint32_t)min(kTilePts, k_end - k0);
    unsigned char* st = stage_raw + (size_t)(t % kStages) * kStageBytes;
    uint64_t* bar = &full_bar[t % kStages];
    ptx::mbar_arrive_expect_tx(bar, np * 29u);
    ptx::bulk_g2s(st + kOffIdx, gIdx + k0, np * 8u, bar);
    ptx::bulk_g2s(st + kOffXC, gXCj + k0, np * 16u, bar);
    ptx::bulk_g2s(st + kOffQ, gQ + k0, np * 4u, bar);
    ptx::bulk_g2s(st + kOffVm, gVm + k0, np, bar);
  };
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) ptx::mbar_init(&full_bar[s], 1);
    ptx::mbar_fence_init();
    for (int t = 0; t < kStages && t < ntile; ++t) issue_tile(t);
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
    const float sa = 1.0f / prm.sigma_a, sb = 1.0f / prm.sigma_b;
    c.sa2 = sa * sa;
    c.sb2 = sb * sb;
    c.C_thresh = prm.C_thresh;
    c.Q_thresh = prm.Q_thresh;
  }
  RaysAcc A;
  {
    const float2 z = make_float2(0.f, 0.f);
    A.TTa = A.TTb = A.TUa = A.TUb = A.UU = A.N6a = A.N6b = z;
#pragma unroll
    for (int p = 0; p < 3; ++p)
      A.TMa[p] = A.TMb[p] = A.UMa[p] = A.UMb[p] = A.MMa[p] = A.MMb[p] = z;
    A.n66 = 0.f;
  }
  __syncthreads();  // barriers initialised before anyone waits on them

  auto stage_of = [&](int t) { return stage_raw + (size_t)(t % kStages) * kStageBytes; };
  auto gbuf_of = [&](int t) {
    return reinterpret_cast<float4*>(stage_raw + (size_t)kStages * kStageBytes +
                                     (size_t)(t & 1) * kGatherBytes);
  };
  auto tile_np = [&](int t) { return min(kTilePts, k_end - (k_begin + t * kTilePts)); };
  // frame-i gather of this thread's points of tile t (stage t must be full): async copies
  auto issue_gather = [&](int t) {
    const unsigned char* st = stage_of(t);
    const int np = tile_np(t);
    float4* gb = gbuf_of(t);
#pragma unroll
    for (int s = 0; s < kPtsPerThread; ++s) {
      const int p = s * kThreads + tid;
      if (p < np) {
        long long ind = reinterpret_cast<const long long*>(st + kOffIdx)[p];
        if (st[kOffVm + p] == 0) ind = 0;  // gn_kernels.cu:677
        ptx::cp_async_16(gb + p, XCi + ind);
      }
    }
    ptx::cp_async_commit();
  };

  uint32_t phase_bits = 0;  // bit s = parity to wait for on stage s
  auto wait_stage = [&](int t) {
    const int s = t % kStages;
    ptx::mbar_wait(&full_bar[s], (phase_bits >> s) & 1u);
    phase_bits ^= 1u << s;
  };
  if (ntile > 0) {
    wait_stage(0);
    issue_gather(0);
  }
  for (int t = 0; t < ntile; ++t) {
    if (t + 1 < ntile) {
      wait_stage(t + 1);
      issue_gather(t + 1);
      ptx::cp_async_wait<1>();  // this thread's gathers of tile t have landed
    } else {
      ptx::cp_async_wait<0>();
    }
    const unsigned char* st = stage_of(t);
    const float4* gb = gbuf_of(t);
    const int np = tile_np(t);
#pragma unroll
    for (int s = 0; s < kPtsPerThread; ++s) {
      const int p = s * kThreads + tid;
      if (p < np) {
        const float4 G = gb[p];
        const float4 xj = reinterpret_cast<const float4*>(st + kOffXC)[p];
        const float q = reinterpret_cast<const float*>(st + kOffQ)[p];
        rays_point(c, A, G.x, G.y, G.z, xj.x, xj.y, xj.z, q, G.w, xj.w, st[kOffVm + p] != 0);
      }
    }
    __syncthreads();  // every reader of stage t % kStages is done
    if (tid == 0 && t + kStages < ntile) issue_tile(t + kStages);
  }
  float S[28], v[7];
  rays_unpack(A, S, v);
  edge_epilogue<kThreads>(S, v, Ti, e, chunk, nchunks, nchunks, partials, tickets, edge_blocks);
}

}  // namespace

// Points per accumulate CTA: minimises ceil(waves) * (points + fixed per-CTA cost) over a few
// multiples of the 768-point tile (2 CTAs per SM; the fixed cost ~ 1500 points covers the
// pipeline fill and the reduction epilogue).
int gn_pts_per_cta(int E, int N) {
  const long long cand[] = {3072, 6144, 9216, 12288, 18432, 24576, 36864};
  const double slots = 2.0 * kNumSMs;
  long long best = 3072;
  double best_cost = 1e300;
  for (long long ppc : cand) {
    const long long nch = ((long long)(N > 0 ? N : 1) + ppc - 1) / ppc;
    const double waves = (double)(nch * (long long)(E > 0 ? E : 1)) / slots;
    const double rounds = waves <= 1.0 ? 1.0 : (double)(long long)(waves + 0.999999);
    const long long eff = ppc < N ? ppc : ((N + 767) / 768) * 768;
    const double cost = rounds * ((double)eff + 1500.0);
    if (cost < best_cost) {
      best_cost = cost;
      best = ppc;
    }
  }
  return (int)best;
}

int launch_gn_pack_points(const float* Xs, const float* Cs, int K, int N, float* XC,
                          cudaStream_t s) {
  const size_t total = (size_t)K * (size_t)N;
  if (total == 0) return FRODO_OK;
  const int blocks = (int)((total + 255) / 256 < (size_t)(8 * kNumSMs) ? (total + 255) / 256
                                                                       : (size_t)(8 * kNumSMs));
  pack_points_kernel<<<blocks, 256, 0, s>>>(Xs, Cs, reinterpret_cast<float4*>(XC), total);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_gn_accumulate(const frodo_gn_params& prm, const float* Twc, const float* Xs,
                         const float* Cs, const float* Kmat, const float* XC,
                         const int32_t* ii_row, const int32_t* jj_row, const int64_t* idx,
                         const uint8_t* valid, const float* Q, int N, int E, float* partials,
                         int32_t* tickets, double* edge_blocks, const int32_t* done,
                         cudaStream_t s) {
  if (E == 0) return FRODO_OK;
  const int ppc = gn_pts_per_cta(E, N);
  const int nchunks = (N + ppc - 1) / ppc;
  dim3 grid(nchunks > 0 ? nchunks : 1, E);
  const bool aligned16 =
      ((reinterpret_cast<uintptr_t>(XC) | reinterpret_cast<uintptr_t>(idx) |
        reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(valid)) & 15) == 0;
  if (prm.mode == FRODO_GN_RAYS && XC != nullptr && (N % 16) == 0 && aligned16) {
    static bool attr_set = false;
    const size_t smem = (size_t)kRaysSmemBytes;
    if (!attr_set) {
      FRODO_CUDA_TRY(cudaFuncSetAttribute(gn_accumulate_rays_packed_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_set = true;
    }
    gn_accumulate_rays_packed_kernel<<<grid, kThreads, smem, s>>>(
        prm, Twc, reinterpret_cast<const float4*>(XC), ii_row, jj_row, idx, valid, Q, N, ppc,
        partials, tickets, edge_blocks, done);
    FRODO_CUDA_TRY(cudaGetLastError());
    return FRODO_OK;
  }
  switch (prm.mode) {
    case FRODO_GN_RAYS:
      gn_accumulate_kernel<FRODO_GN_RAYS><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done);
      break;
    case FRODO_GN_POINTS:
      gn_accumulate_kernel<FRODO_GN_POINTS><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done);
      break;
    case FRODO_GN_CALIB:
      gn_accumulate_kernel<FRODO_GN_CALIB><<<grid, kThreads, 0, s>>>(
          prm, Twc, Xs, Cs, Kmat, ii_row, jj_row, idx, valid, Q, N, ppc, partials, tickets,
          edge_blocks, done);
      break;
    default:
      set_error("unknown GN mode %d", prm.mode);
      return FRODO_ERR_ARG;
  }
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
