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
// rays, fast path (N % 4 == 0, 16-byte aligned inputs): every thread streams 4
// consecutive points per step with 128-bit loads, and the weighted outer
// products run on the packed fp32 pipe (FFMA2, two fp32 lanes per issue slot).
// A residual row is x~ = [x0..x5 | e]: 3 register pairs X0,X1,X2 and the scalar
// e; with WX_p = w X_p the row adds
//   diagonal blocks   Dp  += (WX_p.x, WX_p.y) (x) X_p      (2 FFMA2, one lane dup)
//   off-diagonal      Opq += (WX_p.x, WX_p.y) (x) X_q      (2 FFMA2)
//   gradient          Ep  += WX_p * e                      (1 FFMA2)
// The distance row [r0 r1 r2 | nj | e3] has its own small accumulator set.
// Weights: w = huber(|e| sqrt(q)/sigma) q/sigma^2 = min(1, 1.345 rsqrt(t)) wq
// with t = wq e^2, wq = q/sigma^2 (same value, no sqrt(q)).
// ---------------------------------------------------------------------------
struct RaysAcc {
  float2 Da[3], Db[3];   // diagonal blocks: Da = (S[2p][2p], S[2p][2p+1]), Db = (dup, S[2p+1][2p+1])
  float2 Oa[3], Ob[3];   // blocks (0,1),(0,2),(1,2): rows 2p / 2p+1, columns 2q, 2q+1
  float2 E[3];           // v[2p], v[2p+1]
  float2 Ra, Rb, Rab_a, Rab_b, Re_a, Re_b;  // distance row: pairs A=(r0,r1), B=(r2,nj)
  float2 Rda;            // (A.x*A.x, A.x*A.y)
  float2 Rdb;            // (A.y*A.x, A.y*A.y)
};

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 bc(float a) { return make_float2(a, a); }

__device__ __forceinline__ float huber_w(float wq, float e) {
  // min(1, 1.345 / (|e| sqrt(wq))) * wq ; wq = 0 -> 0
  const float t = wq * e * e;
  return fminf(1.0f, 1.345f * rsqrtf(t)) * wq;
}

__device__ __forceinline__ void rays_row(RaysAcc& A, float w, float2 X0, float2 X1, float2 X2,
                                         float e) {
  const float2 W0 = __fmul2_rn(bc(w), X0);
  const float2 W1 = __fmul2_rn(bc(w), X1);
  const float2 W2 = __fmul2_rn(bc(w), X2);
  A.Da[0] = __ffma2_rn(bc(W0.x), X0, A.Da[0]);
  A.Db[0] = __ffma2_rn(bc(W0.y), X0, A.Db[0]);
  A.Da[1] = __ffma2_rn(bc(W1.x), X1, A.Da[1]);
  A.Db[1] = __ffma2_rn(bc(W1.y), X1, A.Db[1]);
  A.Da[2] = __ffma2_rn(bc(W2.x), X2, A.Da[2]);
  A.Db[2] = __ffma2_rn(bc(W2.y), X2, A.Db[2]);
  A.Oa[0] = __ffma2_rn(bc(W0.x), X1, A.Oa[0]);
  A.Ob[0] = __ffma2_rn(bc(W0.y), X1, A.Ob[0]);
  A.Oa[1] = __ffma2_rn(bc(W0.x), X2, A.Oa[1]);
  A.Ob[1] = __ffma2_rn(bc(W0.y), X2, A.Ob[1]);
  A.Oa[2] = __ffma2_rn(bc(W1.x), X2, A.Oa[2]);
  A.Ob[2] = __ffma2_rn(bc(W1.y), X2, A.Ob[2]);
  A.E[0] = __ffma2_rn(W0, bc(e), A.E[0]);
  A.E[1] = __ffma2_rn(W1, bc(e), A.E[1]);
  A.E[2] = __ffma2_rn(W2, bc(e), A.E[2]);
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
  const float a = rsqrtf(n2j);
  const float nj = n2j * a;
  const float r0 = P0 * a, r1 = P1 * a, r2 = P2 * a;
  const float n2i = fmaf(xi0, xi0, fmaf(xi1, xi1, xi2 * xi2));
  const float bi = rsqrtf(n2i);
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
  const float na = -a;
  const float d00 = fmaf(na * r0, r0, a), d11 = fmaf(na * r1, r1, a), d22 = fmaf(na * r2, r2, a);
  const float d01 = na * r0 * r1, d02 = na * r0 * r2, d12 = na * r1 * r2;
  rays_row(A, w0, f2(d00, d01), f2(d02, 0.0f), f2(r2, -r1), e0);
  rays_row(A, w1, f2(d01, d11), f2(d12, -r2), f2(0.0f, r0), e1);
  rays_row(A, w2, f2(d02, d12), f2(d22, r1), f2(-r0, 0.0f), e2);
  // distance row [r0 r1 r2 | nj | e3]
  const float2 XA = f2(r0, r1), XB = f2(r2, nj);
  const float2 WA = __fmul2_rn(bc(w3), XA), WB = __fmul2_rn(bc(w3), XB);
  A.Rda = __ffma2_rn(bc(WA.x), XA, A.Rda);
  A.Rdb = __ffma2_rn(bc(WA.y), XA, A.Rdb);
  A.Ra = __ffma2_rn(bc(WB.x), XB, A.Ra);
  A.Rb = __ffma2_rn(bc(WB.y), XB, A.Rb);
  A.Rab_a = __ffma2_rn(bc(WA.x), XB, A.Rab_a);
  A.Rab_b = __ffma2_rn(bc(WA.y), XB, A.Rab_b);
  A.Re_a = __ffma2_rn(WA, bc(e3), A.Re_a);
  A.Re_b = __ffma2_rn(WB, bc(e3), A.Re_b);
}

__host__ __device__ constexpr int tri(int n, int m) { return n * 7 - n * (n - 1) / 2 + (m - n); }

// Folds the packed accumulators into the canonical 28 upper-triangular entries + 7.
__device__ __forceinline__ void rays_unpack(const RaysAcc& A, float (&S)[28], float (&v)[7]) {
#pragma unroll
  for (int l = 0; l < 28; ++l) S[l] = 0.f;
  S[tri(0, 0)] = A.Da[0].x + A.Rda.x;
  S[tri(0, 1)] = A.Da[0].y + A.Rda.y;
  S[tri(1, 1)] = A.Db[0].y + A.Rdb.y;
  S[tri(2, 2)] = A.Da[1].x + A.Ra.x;
  S[tri(2, 3)] = A.Da[1].y;
  S[tri(3, 3)] = A.Db[1].y;
  S[tri(4, 4)] = A.Da[2].x;
  S[tri(4, 5)] = A.Da[2].y;
  S[tri(5, 5)] = A.Db[2].y;
  S[tri(0, 2)] = A.Oa[0].x + A.Rab_a.x;
  S[tri(0, 3)] = A.Oa[0].y;
  S[tri(1, 2)] = A.Ob[0].x + A.Rab_b.x;
  S[tri(1, 3)] = A.Ob[0].y;
  S[tri(0, 4)] = A.Oa[1].x;
  S[tri(0, 5)] = A.Oa[1].y;
  S[tri(1, 4)] = A.Ob[1].x;
  S[tri(1, 5)] = A.Ob[1].y;
  S[tri(2, 4)] = A.Oa[2].x;
  S[tri(2, 5)] = A.Oa[2].y;
  S[tri(3, 4)] = A.Ob[2].x;
  S[tri(3, 5)] = A.Ob[2].y;
  S[tri(0, 6)] = A.Rab_a.y;
  S[tri(1, 6)] = A.Rab_b.y;
  S[tri(2, 6)] = A.Ra.y;
  S[tri(6, 6)] = A.Rb.y;
  v[0] = A.E[0].x + A.Re_a.x;
  v[1] = A.E[0].y + A.Re_a.y;
  v[2] = A.E[1].x + A.Re_b.x;
  v[3] = A.E[1].y;
  v[4] = A.E[2].x;
  v[5] = A.E[2].y;
  v[6] = A.Re_b.y;
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
__device__ __forceinline__ void edge_epilogue(float (&S)[28], float (&v)[7], const Sim3f& Ti,
                                              int e, int chunk, int nchunks,
                                              float* __restrict__ partials,
                                              int32_t* __restrict__ tickets,
                                              double* __restrict__ edge_blocks) {
  __shared__ float warp_part[kThreads / 32][kGnBlock];
  __shared__ int is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int l = 0; l < 28; ++l) {
    float x = S[l];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) warp_part[warp][l] = x;
  }
#pragma unroll
  for (int l = 0; l < 7; ++l) {
    float x = v[l];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) warp_part[warp][28 + l] = x;
  }
  __syncthreads();
  float* __restrict__ my_part = partials + ((size_t)e * nchunks + chunk) * kGnBlock;
  if (threadIdx.x < kGnBlock) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) x += warp_part[w][threadIdx.x];
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
    const float* p = partials + (size_t)e * nchunks * kGnBlock + threadIdx.x;
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

  edge_epilogue(S, v, Ti, e, chunk, nchunks, partials, tickets, edge_blocks);
}

// rays fast path: see the comment above RaysAcc.  Requires N % 4 == 0, pts_per_cta % 1024 == 0
// and 16-byte aligned Xs, Cs, idx, Q (valid_match 4-byte aligned).
__global__ void __launch_bounds__(kThreads, 2)
gn_accumulate_rays_v2_kernel(frodo_gn_params prm, const float* __restrict__ Twc,
                             const float* __restrict__ Xs, const float* __restrict__ Cs,
                             const int32_t* __restrict__ ii_row,
                             const int32_t* __restrict__ jj_row, const int64_t* __restrict__ idx,
                             const uint8_t* __restrict__ valid_match, const float* __restrict__ Q,
                             int N, int pts_per_cta, float* __restrict__ partials,
                             int32_t* __restrict__ tickets, double* __restrict__ edge_blocks,
                             const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  const int e = blockIdx.y;
  const int chunk = blockIdx.x;
  const int nchunks = gridDim.x;
  const int ix = ii_row[e], jx = jj_row[e];
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
#pragma unroll
    for (int p = 0; p < 3; ++p) A.Da[p] = A.Db[p] = A.Oa[p] = A.Ob[p] = A.E[p] = z;
    A.Ra = A.Rb = A.Rab_a = A.Rab_b = A.Re_a = A.Re_b = A.Rda = A.Rdb = z;
  }
  const float* __restrict__ Xi_base = Xs + (size_t)ix * N * 3;
  const float* __restrict__ Ci_base = Cs + (size_t)ix * N;
  const float4* __restrict__ Xj4 = reinterpret_cast<const float4*>(Xs + (size_t)jx * N * 3);
  const float4* __restrict__ Cj4 = reinterpret_cast<const float4*>(Cs + (size_t)jx * N);
  const longlong2* __restrict__ idx2 = reinterpret_cast<const longlong2*>(idx + (size_t)e * N);
  const uint32_t* __restrict__ vm4 = reinterpret_cast<const uint32_t*>(valid_match + (size_t)e * N);
  const float4* __restrict__ Q4 = reinterpret_cast<const float4*>(Q + (size_t)e * N);

  const int k_begin = chunk * pts_per_cta;
  const int k_end = min(N, k_begin + pts_per_cta);
  for (int k = k_begin + threadIdx.x * 4; k < k_end; k += kThreads * 4) {
    const int g4 = k >> 2;  // index of this group of 4 points
    // streamed, read-once per-edge data (evict-first) and frame-j points
    const uint32_t vmw = __ldcs(vm4 + g4);
    const longlong2 ia = __ldcs(idx2 + 2 * g4), ib = __ldcs(idx2 + 2 * g4 + 1);
    const float4 q4 = __ldcs(Q4 + g4);
    const float4 cj4 = __ldg(Cj4 + g4);
    const float4 xa = __ldg(Xj4 + 3 * g4), xb = __ldg(Xj4 + 3 * g4 + 1),
                 xc = __ldg(Xj4 + 3 * g4 + 2);
    long long ind[4] = {ia.x, ia.y, ib.x, ib.y};
    bool vm[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      vm[r] = ((vmw >> (8 * r)) & 0xffu) != 0;
      if (!vm[r]) ind[r] = 0;  // gn_kernels.cu:677
    }
    // gathers from frame i
    float xi[4][3], ci[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const float* px = Xi_base + (size_t)ind[r] * 3;
      xi[r][0] = __ldg(px);
      xi[r][1] = __ldg(px + 1);
      xi[r][2] = __ldg(px + 2);
      ci[r] = __ldg(Ci_base + ind[r]);
    }
    rays_point(c, A, xi[0][0], xi[0][1], xi[0][2], xa.x, xa.y, xa.z, q4.x, ci[0], cj4.x, vm[0]);
    rays_point(c, A, xi[1][0], xi[1][1], xi[1][2], xa.w, xb.x, xb.y, q4.y, ci[1], cj4.y, vm[1]);
    rays_point(c, A, xi[2][0], xi[2][1], xi[2][2], xb.z, xb.w, xc.x, q4.z, ci[2], cj4.z, vm[2]);
    rays_point(c, A, xi[3][0], xi[3][1], xi[3][2], xc.y, xc.z, xc.w, q4.w, ci[3], cj4.w, vm[3]);
  }
  float S[28], v[7];
  rays_unpack(A, S, v);
  edge_epilogue(S, v, Ti, e, chunk, nchunks, partials, tickets, edge_blocks);
}

}  // namespace

int gn_pts_per_cta(int E, int N) {
  long long want = ((long long)N * (long long)(E > 0 ? E : 1)) / (kNumSMs * 16);
  long long p = ((want + 1023) / 1024) * 1024;
  if (p < 1024) p = 1024;
  if (p > 8192) p = 8192;
  return (int)p;
}

int launch_gn_accumulate(const frodo_gn_params& prm, const float* Twc, const float* Xs,
                         const float* Cs, const float* Kmat, const int32_t* ii_row,
                         const int32_t* jj_row, const int64_t* idx, const uint8_t* valid,
                         const float* Q, int N, int E, float* partials, int32_t* tickets,
                         double* edge_blocks, const int32_t* done, cudaStream_t s) {
  if (E == 0) return FRODO_OK;
  const int ppc = gn_pts_per_cta(E, N);
  const int nchunks = (N + ppc - 1) / ppc;
  dim3 grid(nchunks > 0 ? nchunks : 1, E);
  const bool aligned16 = ((reinterpret_cast<uintptr_t>(Xs) | reinterpret_cast<uintptr_t>(Cs) |
                           reinterpret_cast<uintptr_t>(idx) | reinterpret_cast<uintptr_t>(Q)) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(valid) & 3) == 0;
  if (prm.mode == FRODO_GN_RAYS && (N % 4) == 0 && aligned16) {
    gn_accumulate_rays_v2_kernel<<<grid, kThreads, 0, s>>>(prm, Twc, Xs, Cs, ii_row, jj_row, idx,
                                                          valid, Q, N, ppc, partials, tickets,
                                                          edge_blocks, done);
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
