// Frame-to-keyframe pose solve of the tracker:
// FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 / solve, mast3r_slam/tracker.py:397-517,
// with geometry.py:17-34 (point_to_ray_dist), :45-52 (act_Sim3), :63-104 (project_calib) and
// nonlinear_optimizer.py:4-33 (check_convergence, huber).
//
// The reference runs up to 50 iterations of ~15 PyTorch kernels each and returns to the host
// every iteration (`.item()` for the cost, the convergence test in Python).  One 512x384 frame
// is 196,608 matched points x 32 bytes = 6.3 MB of iteration-invariant data and a 7x7 system:
// here ONE persistent cooperative kernel (one CTA per SM) reads the inputs from HBM once,
// keeps every CTA's slice of the points resident in shared memory as 8-float records, and
// iterates on chip:
//   per iteration   accumulate the CTA's slice (fp32, per thread) -> warp / CTA tree ->
//                   one fp64 partial [36] per CTA in global memory -> ONE grid barrier ->
//                   every CTA sums all partials in the same fixed order (fp64), solves the 7x7
//                   system, retracts the pose and evaluates the convergence test redundantly,
//                   so all CTAs hold bit-identical poses and take the same decision without a
//                   second barrier (the partials are double-buffered by iteration parity).
// No host synchronisation, no allocation; the only global traffic after the prologue is
// 36 doubles per CTA per iteration through L2.
//
// Residual signs: the reference forms r = z - h(x), J = -dh/dx, g = -A^T b, tau = H^-1 g;
// we accumulate v = sum w e x with e = h(x) - z and x = dh/dx, so tau = -H^-1 v (identical).
#include "common.cuh"
#include "gn_pose_update.cuh"

namespace frodo {

namespace {

constexpr int kTrkThreads = 512;
constexpr int kTrkWarps = kTrkThreads / 32;
constexpr int kRec = 8;        // floats per point record
constexpr int kSums = 36;      // 28 (upper triangle of H) + 7 (v) + 1 (cost)
constexpr int kGroups = 14;    // partial-summation groups (14 * 36 = 504 threads)
constexpr int kSumUnroll = 12; // partials per thread and pass, all loads issued before the adds
constexpr int kMaxSmemPts = 6400;  // 6400 * 32 B = 200 KB of records per CTA

struct TrackArgs {
  const float* Xf;
  const int64_t* idx;
  const float* Xk;
  const float* Qk;
  const uint8_t* valid;
  const float* meas;
  const uint8_t* valid_meas;
  const float* Kmat;       // [3,3] device (calib)
  const float* T_WCf;
  const float* T_WCk;
  float* T_WCf_out;
  float* T_CkCf_out;
  int32_t* info;
  double* costs;
  double* partials;        // [2][grid][36]
  unsigned int* bar;       // grid barrier counter (zeroed before launch)
  float* overflow;         // [8][N] records of the points that do not fit in shared memory
  int N, per, cap;
  int mode, max_iters;
  int raw_calib, width;    // calib: build the measurement from the raw pointmaps in the prologue
  float sa_inv, sb_inv, huber_k, rel_error, delta_norm;
  float fx, fy, cx, cy, ulo, uhi, vlo, vhi, z_eps;
};

__device__ __forceinline__ float huber_w(float r, float k) {
  const float a = fabsf(r);
  return a < k ? 1.0f : __fdividef(k, a);
}

// S (28 upper-triangular entries, row-major) += w x x^T, v += (w e) x over the entries in MASK.
template <unsigned MASK>
__device__ __forceinline__ void add_row(float (&S)[28], float (&v)[7], float& cost,
                                        const float (&x)[7], float w, float e) {
  const float we = w * e;
  cost = fmaf(we, e, cost);
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

// One point.  rec = (xf[3], rk[3], dk, sq) for rays, (xf[3], u, v, log z, sq, -) for calib.
template <int MODE>
__device__ __forceinline__ void point(const TrackArgs& a, const float* T, const float (&rec)[kRec],
                                      float (&S)[28], float (&v)[7], float& cost) {
  float P[3];
  pose_update::act_so3(T + 3, rec, P);
  P[0] = fmaf(P[0], T[7], T[0]);
  P[1] = fmaf(P[1], T[7], T[1]);
  P[2] = fmaf(P[2], T[7], T[2]);
  float x[7];
  if (MODE == FRODO_GN_RAYS) {
    const float sq = rec[7];
    const float n2 = P[0] * P[0] + P[1] * P[1] + P[2] * P[2];
    const float d_inv = rsqrtf(n2);
    const float d = n2 * d_inv;
    const float r0 = P[0] * d_inv, r1 = P[1] * d_inv, r2 = P[2] * d_inv;
    const float e0 = r0 - rec[3], e1 = r1 - rec[4], e2 = r2 - rec[5], e3 = d - rec[6];
    const float sw_r = a.sa_inv * sq, sw_d = a.sb_inv * sq;
    const float wr = sw_r * sw_r, wd = sw_d * sw_d;
    const float w0 = huber_w(sw_r * e0, a.huber_k) * wr;
    const float w1 = huber_w(sw_r * e1, a.huber_k) * wr;
    const float w2 = huber_w(sw_r * e2, a.huber_k) * wr;
    const float w3 = huber_w(sw_d * e3, a.huber_k) * wd;
    // d(r,d)/dP [I, -[P]x, P]: (I - r r^T)/d for the ray rows (their [P]x part is -[r]x and
    // their scale column vanishes), r^T for the distance row (geometry.py:26-33, :49-52)
    const float dxx = fmaf(-r0, r0, 1.0f) * d_inv;
    const float dyy = fmaf(-r1, r1, 1.0f) * d_inv;
    const float dzz = fmaf(-r2, r2, 1.0f) * d_inv;
    const float dxy = -r0 * r1 * d_inv;
    const float dxz = -r0 * r2 * d_inv;
    const float dyz = -r1 * r2 * d_inv;
    x[0] = dxx; x[1] = dxy; x[2] = dxz; x[3] = 0.f; x[4] = r2; x[5] = -r1; x[6] = 0.f;
    add_row<0x37u>(S, v, cost, x, w0, e0);
    x[0] = dxy; x[1] = dyy; x[2] = dyz; x[3] = -r2; x[4] = 0.f; x[5] = r0; x[6] = 0.f;
    add_row<0x2Fu>(S, v, cost, x, w1, e1);
    x[0] = dxz; x[1] = dyz; x[2] = dzz; x[3] = r1; x[4] = -r0; x[5] = 0.f; x[6] = 0.f;
    add_row<0x1Fu>(S, v, cost, x, w2, e2);
    x[0] = r0; x[1] = r1; x[2] = r2; x[3] = 0.f; x[4] = 0.f; x[5] = 0.f; x[6] = d;
    add_row<0x47u>(S, v, cost, x, w3, e3);
  } else {  // calib: project_calib, geometry.py:63-104
    const bool valid_z = P[2] > a.z_eps;
    const float zinv = valid_z ? __fdividef(1.0f, P[2]) : 0.0f;
    const float logz = valid_z ? logf(P[2]) : 0.0f;
    const float xz = P[0] * zinv, yz = P[1] * zinv;
    const float u = fmaf(a.fx, xz, a.cx);
    const float vv = fmaf(a.fy, yz, a.cy);
    const bool ok = valid_z && (u > a.ulo) && (u < a.uhi) && (vv > a.vlo) && (vv < a.vhi);
    const float sq = ok ? rec[6] : 0.0f;
    const float e0 = u - rec[3], e1 = vv - rec[4], e2 = logz - rec[5];
    const float sw_p = a.sa_inv * sq, sw_d = a.sb_inv * sq;
    const float wp = sw_p * sw_p, wd = sw_d * sw_d;
    const float w0 = huber_w(sw_p * e0, a.huber_k) * wp;
    const float w1 = huber_w(sw_p * e1, a.huber_k) * wp;
    const float w2 = huber_w(sw_d * e2, a.huber_k) * wd;
    x[0] = a.fx * zinv; x[1] = 0.f; x[2] = -a.fx * xz * zinv; x[3] = -a.fx * xz * yz;
    x[4] = a.fx * fmaf(xz, xz, 1.0f); x[5] = -a.fx * yz; x[6] = 0.f;
    add_row<0x3Du>(S, v, cost, x, w0, e0);
    x[0] = 0.f; x[1] = a.fy * zinv; x[2] = -a.fy * yz * zinv; x[3] = -a.fy * fmaf(yz, yz, 1.0f);
    x[4] = a.fy * xz * yz; x[5] = a.fy * xz; x[6] = 0.f;
    add_row<0x3Eu>(S, v, cost, x, w1, e1);
    x[0] = 0.f; x[1] = 0.f; x[2] = zinv; x[3] = yz; x[4] = -xz; x[5] = 0.f; x[6] = 1.f;
    add_row<0x5Cu>(S, v, cost, x, w2, e2);
  }
}

// T_CkCf = T_WCk^-1 * T_WCf (tracker.py:427, :475), fp32.
__device__ void rel_pose(const float* Tk, const float* Tf, float* R) {
  const float si = 1.0f / Tk[7];
  const float qi[4] = {-Tk[3], -Tk[4], -Tk[5], Tk[6]};
  pose_update::quat_comp(qi, Tf + 3, R + 3);
  const float d[3] = {Tf[0] - Tk[0], Tf[1] - Tk[1], Tf[2] - Tk[2]};
  pose_update::act_so3(qi, d, R);
  R[0] *= si; R[1] *= si; R[2] *= si;
  R[7] = si * Tf[7];
}

// T_WCf = T_WCk * T_CkCf (tracker.py:463, :515).
__device__ void compose_pose(const float* A, const float* B, float* C) {
  float t[3];
  pose_update::act_so3(A + 3, B, t);
  C[0] = fmaf(t[0], A[7], A[0]);
  C[1] = fmaf(t[1], A[7], A[1]);
  C[2] = fmaf(t[2], A[7], A[2]);
  pose_update::quat_comp(A + 3, B + 3, C + 3);
  C[7] = A[7] * B[7];
}

// 7x7 fp64 Cholesky solve of H tau = -v from the 36 sums; false if H is not positive definite
// (torch.linalg.cholesky raises in the reference, tracker.py:407; track() reports failure).
__device__ __forceinline__ bool solve7(const double* sums, double* tau) {
  double L[7][7];  // fully unrolled below: lives in registers
  {
    int l = 0;
#pragma unroll
    for (int n = 0; n < 7; ++n)
#pragma unroll
      for (int m = n; m < 7; ++m) L[m][n] = sums[l++];
  }
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    double d = L[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d = fma(-L[j][k], L[j][k], d);
    ok = ok && (d > 0.0) && isfinite(d);
    const double inv = rsqrt(d);
    L[j][j] = inv;  // the diagonal holds 1 / l_jj
#pragma unroll
    for (int i = j + 1; i < 7; ++i) {
      double t = L[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) t = fma(-L[i][k], L[j][k], t);
      L[i][j] = t * inv;
    }
  }
  if (!ok) return false;
  double y[7];
#pragma unroll
  for (int i = 0; i < 7; ++i) {
    double t = -sums[28 + i];
#pragma unroll
    for (int k = 0; k < i; ++k) t = fma(-L[i][k], y[k], t);
    y[i] = t * L[i][i];
  }
#pragma unroll
  for (int i = 6; i >= 0; --i) {
    double t = y[i];
#pragma unroll
    for (int k = i + 1; k < 7; ++k) t = fma(-L[k][i], tau[k], t);
    tau[i] = t * L[i][i];
  }
  return true;
}

__device__ __forceinline__ unsigned int ld_acquire(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int MODE>
__global__ void __launch_bounds__(kTrkThreads, 1) track_pose_kernel(const TrackArgs a_in) {
  TrackArgs a = a_in;
  if (MODE == FRODO_GN_CALIB) {  // decompose_K, geometry.py:55-60
    a.fx = __ldg(a.Kmat + 0); a.fy = __ldg(a.Kmat + 4);
    a.cx = __ldg(a.Kmat + 2); a.cy = __ldg(a.Kmat + 5);
  }
  extern __shared__ __align__(16) float rec_s[];  // [kRec][cap]
  __shared__ float warp_part[kTrkWarps][kSums];
  __shared__ double grp_part[kGroups][kSums];
  __shared__ double sums[kSums];
  __shared__ float T_s[8];
  __shared__ int ctl_s[2];  // 0: stop, 1: status

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = blockIdx.x, G = gridDim.x;
  const int start = cta * a.per;
  const int cnt = max(0, min(a.per, a.N - start));
  const int cap = a.cap;

  // ---- prologue: gather the frame's points, form the iteration-invariant records ---------
  // Four points per thread and pass: all streamed loads (idx, valid, Q, Xk | meas) are issued
  // first, then the four data-dependent gathers of Xf, then the arithmetic -- the prologue is
  // latency-bound (<= 3 points per thread at 512x384), so loads in flight are what count.
  constexpr int kPro = 4;
  for (int p0 = tid; p0 < cnt; p0 += kPro * kTrkThreads) {
    long long m[kPro];
    float q[kPro], km[kPro][3], xf[kPro][3];
    bool ok[kPro];
#pragma unroll
    for (int u = 0; u < kPro; ++u) {
      const int p = p0 + u * kTrkThreads;
      const int n = start + min(p, cnt - 1);  // clamped: loads stay unconditional
      m[u] = a.idx ? a.idx[n] : (long long)n;
      ok[u] = a.valid[n] != 0;
      q[u] = a.Qk[n];
      const bool raw = MODE != FRODO_GN_RAYS && a.raw_calib != 0;
      const float* src = ((MODE == FRODO_GN_RAYS || raw) ? a.Xk : a.meas) + (size_t)n * 3;
      km[u][0] = src[0]; km[u][1] = src[1]; km[u][2] = src[2];
      if (MODE != FRODO_GN_RAYS && !raw) ok[u] = ok[u] && a.valid_meas[n] != 0;
    }
#pragma unroll
    for (int u = 0; u < kPro; ++u) {
      const float* x = a.Xf + m[u] * 3;
      xf[u][0] = x[0]; xf[u][1] = x[1]; xf[u][2] = x[2];
    }
#pragma unroll
    for (int u = 0; u < kPro; ++u) {
      const int p = p0 + u * kTrkThreads;
      if (p >= cnt) break;
      const int n = start + p;
      float r[kRec];
      r[0] = xf[u][0]; r[1] = xf[u][1]; r[2] = xf[u][2];
      if (MODE != FRODO_GN_RAYS && a.raw_calib != 0) {
        // the calibrated branch of get_points_poses (tracker.py:383-393): both points are put
        // back on their own pixel ray at their depth (constrain_points_to_ray / backproject,
        // geometry.py:37-42,107-116), the keyframe's measurement is (u, v, log z)
        const int mf = (int)m[u];
        const float uf = (float)(mf % a.width), vf = (float)(mf / a.width);
        const float zf = xf[u][2];
        r[0] = zf * __fdiv_rn(uf - a.cx, a.fx);
        r[1] = zf * __fdiv_rn(vf - a.cy, a.fy);
        r[2] = zf * 1.0f;
        const float zk = km[u][2];
        const bool vz = zk > a.z_eps;
        ok[u] = ok[u] && vz;
        km[u][0] = (float)(n % a.width);
        km[u][1] = (float)(n / a.width);
        km[u][2] = vz ? logf(zk) : 0.0f;
      }
      if (MODE == FRODO_GN_RAYS) {
        const float k0 = km[u][0], k1 = km[u][1], k2 = km[u][2];
        const float d = sqrtf(k0 * k0 + k1 * k1 + k2 * k2);
        const float d_inv = 1.0f / d;
        r[3] = d_inv * k0; r[4] = d_inv * k1; r[5] = d_inv * k2; r[6] = d;
        r[7] = ok[u] ? sqrtf(q[u]) : 0.0f;
      } else {
        r[3] = km[u][0]; r[4] = km[u][1]; r[5] = km[u][2];
        r[6] = ok[u] ? sqrtf(q[u]) : 0.0f;
        r[7] = 0.0f;
      }
      if (!ok[u]) {  // zero information: keep the arithmetic finite whatever the inputs hold
        r[0] = 0.f; r[1] = 0.f; r[2] = 1.f; r[3] = 0.f; r[4] = 0.f;
        r[5] = MODE == FRODO_GN_RAYS ? 1.f : 0.f;
        r[6] = MODE == FRODO_GN_RAYS ? 1.f : 0.f;
        r[7] = 0.f;
      }
      if (p < cap) {
#pragma unroll
        for (int c = 0; c < kRec; ++c) rec_s[c * cap + p] = r[c];
      } else {
#pragma unroll
        for (int c = 0; c < kRec; ++c) a.overflow[(size_t)c * a.N + n] = r[c];
      }
    }
  }
  if (tid == 0) {
    float Tk[8], Tf[8], R[8];
    for (int c = 0; c < 8; ++c) { Tk[c] = a.T_WCk[c]; Tf[c] = a.T_WCf[c]; }
    rel_pose(Tk, Tf, R);
    for (int c = 0; c < 8; ++c) T_s[c] = R[c];
    ctl_s[0] = 0;
    ctl_s[1] = 0;
  }
  __syncthreads();

  double old_cost = 0.0;  // thread 0 only
  bool have_old = false;
  int iters = 0;
  for (int it = 0; it < a.max_iters; ++it) {
    float T[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) T[c] = T_s[c];
    float S[28], v[7], cost = 0.0f;
#pragma unroll
    for (int l = 0; l < 28; ++l) S[l] = 0.0f;
#pragma unroll
    for (int l = 0; l < 7; ++l) v[l] = 0.0f;
    for (int p = tid; p < cnt; p += kTrkThreads) {
      float r[kRec];
      if (p < cap) {
#pragma unroll
        for (int c = 0; c < kRec; ++c) r[c] = rec_s[c * cap + p];
      } else {
#pragma unroll
        for (int c = 0; c < kRec; ++c) r[c] = a.overflow[(size_t)c * a.N + start + p];
      }
      point<MODE>(a, T, r, S, v, cost);
    }
    // ---- CTA reduction: warp reduce-scatter (lane m ends with the warp total of entry m,
    // 31 exchanges for 32 entries) + butterflies for the last 4, then fp64 across the warps --
    {
      float r32[32];
#pragma unroll
      for (int l = 0; l < 28; ++l) r32[l] = S[l];
#pragma unroll
      for (int l = 0; l < 4; ++l) r32[28 + l] = v[l];
#pragma unroll
      for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
          const float send = up ? r32[i] : r32[i + half];
          const float keep = up ? r32[i + half] : r32[i];
          r32[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
      }
      warp_part[warp][lane] = r32[0];
#pragma unroll
      for (int l = 32; l < kSums; ++l) {
        float x = l < 35 ? v[l - 28] : cost;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) warp_part[warp][l] = x;
      }
    }
    __syncthreads();
    double* part = a.partials + (size_t)(it & 1) * G * kSums;
    if (tid < kSums) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < kTrkWarps; ++w) s += (double)warp_part[w][tid];
      part[(size_t)cta * kSums + tid] = s;
    }
    __syncthreads();
    // ---- grid barrier (all CTAs are co-resident: cooperative launch) ---------------------
    if (tid == 0) {
      __threadfence();
      atomicAdd(a.bar, 1u);
      const unsigned int target = (unsigned int)(it + 1) * (unsigned int)G;
      while (ld_acquire(a.bar) < target) {
      }
    }
    __syncthreads();
    // ---- every CTA: same fixed-order fp64 sum of all partials, same 7x7 solve -------------
    if (tid < kGroups * kSums) {
      const int grp = tid / kSums, l = tid - grp * kSums;
      double s = 0.0;
      for (int c0 = grp; c0 < G; c0 += kGroups * kSumUnroll) {
        double vals[kSumUnroll];
#pragma unroll
        for (int u = 0; u < kSumUnroll; ++u) {
          const int c = c0 + u * kGroups;
          vals[u] = c < G ? __ldcg(part + (size_t)c * kSums + l) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < kSumUnroll; ++u) s += vals[u];
      }
      grp_part[grp][l] = s;
    }
    __syncthreads();
    if (tid < kSums) {
      double s = 0.0;
#pragma unroll
      for (int g = 0; g < kGroups; ++g) s += grp_part[g][tid];
      sums[tid] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double tau[7];
      const double new_cost = 0.5 * sums[35];
      if (!solve7(sums, tau)) {
        ctl_s[0] = 1;
        ctl_s[1] = 1;
      } else {
        float xi[7], nrm = 0.0f;
#pragma unroll
        for (int c = 0; c < 7; ++c) {
          xi[c] = (float)tau[c];
          nrm = fmaf(xi[c], xi[c], nrm);
        }
        float P[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) P[c] = T_s[c];
        pose_update::retract_pose(P, xi);
#pragma unroll
        for (int c = 0; c < 8; ++c) T_s[c] = P[c];
        iters = it + 1;
        if (cta == 0 && a.costs) a.costs[it] = new_cost;
        // check_convergence, nonlinear_optimizer.py:4-25: on the first step old_cost is inf,
        // rel_dec is nan and only the step norm can stop the loop
        bool conv = sqrtf(nrm) < a.delta_norm;
        if (have_old) conv = conv || fabs((old_cost - new_cost) / old_cost) < (double)a.rel_error;
        old_cost = new_cost;
        have_old = true;
        if (conv) ctl_s[0] = 1;
      }
    }
    __syncthreads();
    if (ctl_s[0]) break;
  }
  if (cta == 0 && tid == 0) {
    float Tk[8], R[8], Tw[8];
    for (int c = 0; c < 8; ++c) { Tk[c] = a.T_WCk[c]; R[c] = T_s[c]; }
    compose_pose(Tk, R, Tw);
    for (int c = 0; c < 8; ++c) { a.T_CkCf_out[c] = R[c]; a.T_WCf_out[c] = Tw[c]; }
    if (a.info) { a.info[0] = iters; a.info[1] = ctl_s[1]; }
  }
}

struct TrackWorkspace {
  unsigned int* bar;
  double* partials;
  float* overflow;
};

size_t track_carve(void* base, int N, int grid_max, TrackWorkspace* ws) {
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  size_t off = 0;
  const size_t o_bar = off; off = up(off + 64);
  const size_t o_part = off; off = up(off + (size_t)2 * grid_max * kSums * sizeof(double));
  const size_t o_over = off; off = up(off + (size_t)kRec * (size_t)(N > 0 ? N : 1) * sizeof(float));
  if (base && ws) {
    char* b = static_cast<char*>(base);
    ws->bar = reinterpret_cast<unsigned int*>(b + o_bar);
    ws->partials = reinterpret_cast<double*>(b + o_part);
    ws->overflow = reinterpret_cast<float*>(b + o_over);
  }
  return off;
}

constexpr int kGridMax = 1024;  // upper bound on the SM count the workspace is sized for

}  // namespace

size_t track_workspace_bytes(int N) { return track_carve(nullptr, N, kGridMax, nullptr); }

int launch_track_pose(const frodo_track_params& prm, const float* Xf, const int64_t* idx,
                      const float* Xk, const float* Qk, const uint8_t* valid, const float* meas,
                      const uint8_t* valid_meas, const float* Kmat, const float* T_WCf,
                      const float* T_WCk, int N, float* T_WCf_out, float* T_CkCf_out,
                      int32_t* info, double* costs, void* workspace, cudaStream_t s) {
  int dev = 0, sms = 0, coop = 0;
  FRODO_CUDA_TRY(cudaGetDevice(&dev));
  FRODO_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  FRODO_CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  if (!coop || sms <= 0 || sms > kGridMax) {
    set_error("track_pose needs cooperative launch (device reports %d, %d SMs)", coop, sms);
    return FRODO_ERR_CUDA;
  }
  TrackWorkspace ws;
  track_carve(workspace, N, kGridMax, &ws);
  TrackArgs a;
  a.Xf = Xf; a.idx = idx; a.Xk = Xk; a.Qk = Qk; a.valid = valid; a.meas = meas;
  a.valid_meas = valid_meas; a.Kmat = Kmat; a.T_WCf = T_WCf; a.T_WCk = T_WCk; a.T_WCf_out = T_WCf_out;
  a.T_CkCf_out = T_CkCf_out; a.info = info; a.costs = costs; a.partials = ws.partials;
  a.bar = ws.bar; a.overflow = ws.overflow;
  int grid = (N + kTrkThreads - 1) / kTrkThreads;
  grid = grid < 1 ? 1 : (grid > sms ? sms : grid);
  a.N = N;
  a.per = (N + grid - 1) / grid;
  if (a.per < 1) a.per = 1;
  a.cap = a.per < kMaxSmemPts ? a.per : kMaxSmemPts;
  a.mode = prm.mode; a.max_iters = prm.max_iters;
  a.raw_calib = prm.raw_calib; a.width = prm.width > 0 ? prm.width : 1;
  a.sa_inv = 1.0f / prm.sigma_a; a.sb_inv = 1.0f / prm.sigma_b;
  a.huber_k = prm.huber; a.rel_error = prm.rel_error; a.delta_norm = prm.delta_norm;
  a.fx = a.fy = 1.0f; a.cx = a.cy = 0.0f;
  a.ulo = (float)prm.pixel_border; a.uhi = (float)(prm.width - 1 - prm.pixel_border);
  a.vlo = (float)prm.pixel_border; a.vhi = (float)(prm.height - 1 - prm.pixel_border);
  a.z_eps = prm.z_eps;
  const size_t smem = (size_t)kRec * a.cap * sizeof(float);
  void* fn = prm.mode == FRODO_GN_RAYS ? (void*)track_pose_kernel<FRODO_GN_RAYS>
                                       : (void*)track_pose_kernel<FRODO_GN_CALIB>;
  FRODO_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FRODO_CUDA_TRY(cudaMemsetAsync(ws.bar, 0, 64, s));
  void* args[] = {(void*)&a};
  FRODO_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kTrkThreads), args, smem, s));
  return FRODO_OK;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops).  The
// workspace of the solve above lives in a file-scope device array sized for frames of up to
// kStaticMaxPoints points; every GPU that loads the library holds these bytes until exit.
// All calls on a device share the array: the grid-barrier words are cleared by every call
// (launch_track_pose does that with a stream-ordered memset) and the rest is written before it
// is read, so stale contents of the previous call are harmless -- but two calls that overlap
// on the GPU (different streams) race on it.
namespace {
constexpr int kStaticMaxPoints = 512 * 384;  // one 512x384 frame (BASELINE.json's size)
constexpr size_t static_up(size_t x) { return (x + 255) / 256 * 256; }
// the same layout track_carve() produces for N = kStaticMaxPoints (checked at set-up)
constexpr size_t kStaticBytes = static_up(64) +
                                static_up((size_t)2 * kGridMax * kSums * sizeof(double)) +
                                static_up((size_t)kRec * kStaticMaxPoints * sizeof(float));
static __device__ __align__(256) unsigned char g_track_scratch[kStaticBytes];
}  // namespace

int track_static_max_points() { return kStaticMaxPoints; }

int launch_track_pose_static(const frodo_track_params& prm, const float* Xf, const int64_t* idx,
                             const float* Xk, const float* Qk, const uint8_t* valid,
                             const float* meas, const uint8_t* valid_meas, const float* Kmat,
                             const float* T_WCf, const float* T_WCk, int N, float* T_WCf_out,
                             float* T_CkCf_out, int32_t* info, double* costs, cudaStream_t s) {
  if (N > kStaticMaxPoints || track_workspace_bytes(N) > kStaticBytes) {
    set_error("frodo_track_pose_static: N = %d exceeds the built-in scratch (at most %d points); "
              "use frodo_track_pose with a caller-provided workspace", N, kStaticMaxPoints);
    return FRODO_ERR_ARG;
  }
  // In host code `g_track_scratch` is only a handle for the symbol, not a device address:
  // resolve it once per device, with that device current.
  static void* dev_ptr[kMaxDevices] = {};
  int dev = -1;
  FRODO_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("frodo_track_pose_static: device ordinal %d not supported", dev);
    return FRODO_ERR_ARG;
  }
  if (dev_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    FRODO_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_track_scratch));
    if (ptr == nullptr || (reinterpret_cast<uintptr_t>(ptr) & 255u) != 0) {
      set_error("frodo_track_pose_static: could not resolve the scratch symbol");
      return FRODO_ERR_CUDA;
    }
    dev_ptr[dev] = ptr;
  }
  return launch_track_pose(prm, Xf, idx, Xk, Qk, valid, meas, valid_meas, Kmat, T_WCf, T_WCk, N,
                           T_WCf_out, T_CkCf_out, info, costs, dev_ptr[dev], s);
}

}  // namespace frodo
