// Extra terms of gauss_newton_rays_odom (gn.cpp:29-62, gn_kernels.cu:1792-1934): relative-pose
// (odometry) constraints between keyframes and a per-keyframe scale prior, added to the dense
// fp64 normal equations after the visual terms have been assembled.
//
// Residual of one odometry edge (i -> j, measured relative pose [R_m, t_m]; the reference's
// odom_constraint_kernel_left_perturb_log, :1106-1286):
//   R_err = R_m^T R_i^T R_j,  phi = Log(R_err),
//   rho   = J_l(phi)^-1 (R_m^T (R_i^T (t_j - t_i) / s_j) - R_m^T t_m),   r = [rho, phi],
// weights W = diag(huber(r / sigma) / sigma^2), Jacobian w.r.t. pose j
//   J_j = [ R_j^T   [-R_j^T t_j]x R_j^T   -R_m^T t_pred ]     (6 x 7, tangent [tau, theta, sigma])
//         [ 0       R_j^T                  0            ]
// and J_i = -J_j with its scale column zeroed.  Hence, as for the visual edges (gn_accumulate.cu),
// ONE symmetric 7x7  S = J_j^T W J_j  and one 7-vector  v = J_j^T W r  carry the whole 14x14 block:
//   H_jj = S,   H_ii = S with row/column 6 zeroed,   H_ij = -S with ROW 6 zeroed,
//   g_j = v,    g_i = -v with entry 6 zeroed,
// which is what odom_blocks_kernel stores (28 + 7 doubles per edge) and odom_add_kernel expands.
// S and v are fp32 sums over the six residual rows in row order, so every entry is bit-equal to
// the corresponding entry of the reference's 14x14 product.
//
// Parity note: phi goes through acosf(0.5 (trace - 1)), which quantises angles below ~5e-4 rad;
// one ulp in R_err moves phi by 3e-4 rad at an information of 1e6.  The chain quaternion -> R ->
// R_err therefore uses explicit _rn intrinsics in the operation order nvcc gives the reference
// (read from its SASS; the CPU restatement used by the tests follows it too); everything after Log() is
// well-conditioned and written in this file's own form.
//   * odom_add_kernel: one CTA per lower 7x7 block of H (like assemble_kernel), visiting the
//     odometry edges in order (deterministic fp64 sums), SparseBlock::update_lhs/rhs semantics
//     (:72-122: blocks with a fixed keyframe are skipped), plus the scale prior of
//     apply_scale_prior_cpu (:1716-1790) on the diagonal.
// Reference quirks kept on purpose (parity): odometry ids index Twc rows RAW but are ranked
// against the visual keyframe list for the system (:1121-1122, :1829); the "current scale" of
// the prior is read from column 6 of Twc, which is q_w, not s (:1736, :1762).
#include "gn_internal.cuh"

namespace frodo {

namespace {

constexpr int kOdomBlock = kGnBlock;  // 28 upper-triangular entries of S, then v

struct Rot3 {
  float m[9];  // row-major
  __device__ float operator()(int r, int c) const { return m[3 * r + c]; }
};

// Rotation matrix of a (not renormalised) quaternion xyzw; operation order as in the SASS of the
// reference's quat_to_rot (:562-574): 1 - 2 y^2 - 2 z^2 -> fma(-z, 2z, fma(-y, 2y, 1)) etc.
__device__ __forceinline__ Rot3 rot_from_quat(const float* q) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float x2 = __fadd_rn(x, x), y2 = __fadd_rn(y, y), z2 = __fadd_rn(z, z);
  const float wz = __fmul_rn(w, z2), wy = __fmul_rn(w, y2), wx = __fmul_rn(w, x2);
  const float ax = __fmaf_rn(-x, x2, 1.0f), ay = __fmaf_rn(-y, y2, 1.0f);
  Rot3 R;
  R.m[0] = __fmaf_rn(-z, z2, ay);
  R.m[1] = __fmaf_rn(y, x2, -wz);
  R.m[2] = __fmaf_rn(z, x2, wy);
  R.m[3] = __fmaf_rn(y, x2, wz);
  R.m[4] = __fmaf_rn(-z, z2, ax);
  R.m[5] = __fmaf_rn(z, y2, -wx);
  R.m[6] = __fmaf_rn(z, x2, -wy);
  R.m[7] = __fmaf_rn(z, y2, wx);
  R.m[8] = __fmaf_rn(-y, y2, ax);
  return R;
}

// A^T B without forming the transpose; inner index ascending, one fma per term.
__device__ __forceinline__ Rot3 mul_tn(const Rot3& A, const Rot3& B) {
  Rot3 C;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c)
      C.m[3 * r + c] = __fmaf_rn(A(2, r), B(2, c), __fmaf_rn(A(1, r), B(1, c), __fmul_rn(A(0, r), B(0, c))));
  return C;
}

// A^T x; term order (0, 1 first, then 2) as the reference's 3x3 matrix-vector product compiles.
__device__ __forceinline__ void mul_tv(const Rot3& A, const float* x, float* y) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
    y[r] = __fmaf_rn(A(2, r), x[2], __fmaf_rn(A(0, r), x[0], __fmul_rn(A(1, r), x[1])));
}

__device__ __forceinline__ float huber_weight(float r) {  // :214-217 (fp64 compare and divide)
  const float a = fabsf(r);
  return (double)a < 1.345 ? 1.0f : (float)(1.345 / (double)a);
}

// SO(3) logarithm with the reference's branch thresholds (:1007-1041): zero below 1e-6 rad, the
// symmetric-part formula within 1e-6 of pi (largest diagonal entry first), else the
// antisymmetric part scaled by angle / (2 sin angle).
__device__ void so3_log(const Rot3& R, float* phi) {
  const float c = fmaxf(-1.0f, fminf(1.0f, 0.5f * ((R(0, 0) + R(1, 1) + R(2, 2)) - 1.0f)));
  const float angle = acosf(c);
  if (fabsf(angle) < 1e-6f) {
    phi[0] = phi[1] = phi[2] = 0.0f;
    return;
  }
  if (fabsf((float)((double)angle - 3.14159265358979323846)) < 1e-6f) {
    const int a = (R(0, 0) > R(1, 1) && R(0, 0) > R(2, 2)) ? 0 : (R(1, 1) > R(2, 2) ? 1 : 2);
    const int b = a == 0 ? 1 : 0, d = a == 2 ? 1 : 2;  // the other two axes, ascending
    const float sv = 2.0f * sqrtf(R(a, a) - R(b, b) - R(d, d) + 1.0f);
    phi[a] = 0.5f * sv * angle;
    phi[b] = (R(a, b) + R(b, a)) / sv * angle;
    phi[d] = (R(a, d) + R(d, a)) / sv * angle;
    return;
  }
  const float f = angle / (2.0f * sinf(angle));
  phi[0] = f * (R(2, 1) - R(1, 2));
  phi[1] = f * (R(0, 2) - R(2, 0));
  phi[2] = f * (R(1, 0) - R(0, 1));
}

// y = J_l(phi)^-1 x = x - 0.5 phi x x + k phi x (phi x x),  k = 1/a^2 - (1 + cos a) / (2 a sin a)
// (:1044-1074; k -> 0 term dropped below 1e-6 rad as there).
__device__ void left_jacobian_inv_apply(const float* phi, const float* x, float* y) {
  const float a = sqrtf(__fmaf_rn(phi[2], phi[2], __fmaf_rn(phi[0], phi[0], __fmul_rn(phi[1], phi[1]))));
  const float c1[3] = {phi[1] * x[2] - phi[2] * x[1], phi[2] * x[0] - phi[0] * x[2],
                       phi[0] * x[1] - phi[1] * x[0]};
  float k = 0.0f;
  if (!(fabsf(a) < 1e-6f)) k = (1.0f / (a * a)) - ((1.0f + cosf(a)) / (2.0f * a * sinf(a)));
  const float c2[3] = {phi[1] * c1[2] - phi[2] * c1[1], phi[2] * c1[0] - phi[0] * c1[2],
                       phi[0] * c1[1] - phi[1] * c1[0]};
#pragma unroll
  for (int r = 0; r < 3; ++r) y[r] = x[r] - 0.5f * c1[r] + k * c2[r];
}

__global__ void __launch_bounds__(64)
odom_blocks_kernel(const float* __restrict__ Twc, const float* __restrict__ Tij, int stride,
                   const int64_t* __restrict__ odom_ii, const int64_t* __restrict__ odom_jj,
                   int Eo, float sigma_t, float sigma_r, double* __restrict__ blocks,
                   const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Eo) return;
  const float* Pi = Twc + 8 * odom_ii[idx];
  const float* Pj = Twc + 8 * odom_jj[idx];
  const float* M = Tij + (size_t)stride * idx;
  const float ti[3] = {Pi[0], Pi[1], Pi[2]}, tj[3] = {Pj[0], Pj[1], Pj[2]};
  const float tm[3] = {M[0], M[1], M[2]};
  const float qi[4] = {Pi[3], Pi[4], Pi[5], Pi[6]}, qj[4] = {Pj[3], Pj[4], Pj[5], Pj[6]};
  const float qm[4] = {M[3], M[4], M[5], M[6]};
  const float sj = Pj[7];
  const Rot3 Rm = rot_from_quat(qm), Ri = rot_from_quat(qi), Rj = rot_from_quat(qj);

  // ---- residual r = [rho, phi] --------------------------------------------------------------
  const Rot3 Rerr = mul_tn(Rm, mul_tn(Ri, Rj));
  const float dt[3] = {__fsub_rn(tj[0], ti[0]), __fsub_rn(tj[1], ti[1]), __fsub_rn(tj[2], ti[2])};
  float t_pred[3], rm_tp[3], rm_tm[3];
  mul_tv(Ri, dt, t_pred);
  const float sj_inv = (sj == 0.0f) ? 1e6f : __fdiv_rn(1.0f, sj);  // :1149 guards a zero scale
#pragma unroll
  for (int k = 0; k < 3; ++k) t_pred[k] = __fmul_rn(sj_inv, t_pred[k]);
  mul_tv(Rm, t_pred, rm_tp);  // also d rho / d sigma_j, up to sign
  mul_tv(Rm, tm, rm_tm);
  const float t_err[3] = {__fsub_rn(rm_tp[0], rm_tm[0]), __fsub_rn(rm_tp[1], rm_tm[1]),
                          __fsub_rn(rm_tp[2], rm_tm[2])};
  float r[6];
  so3_log(Rerr, r + 3);
  left_jacobian_inv_apply(r + 3, t_err, r);
  float w[6];
  const float info_t = 1.0f / (sigma_t * sigma_t), info_r = 1.0f / (sigma_r * sigma_r);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    w[k] = info_t * huber_weight(r[k] / sigma_t);
    w[3 + k] = info_r * huber_weight(r[3 + k] / sigma_r);
  }

  // ---- J_j (6 x 7), rows 0-2: [Rj^T | [c]x Rj^T | -Rm^T t_pred], rows 3-5: [0 | Rj^T | 0] ------
  float cpos[3];  // c = -Rj^T t_j: camera j's origin seen from frame j
  mul_tv(Rj, tj, cpos);
#pragma unroll
  for (int k = 0; k < 3; ++k) cpos[k] = -cpos[k];
  float J[6][7];
#pragma unroll
  for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float rjt = Rj(c, rr);  // (Rj^T)[rr][c]
      J[rr][c] = rjt;
      // row rr of [c]x Rj^T = c[r1] * (Rj^T)[r2][.] - c[r2] * (Rj^T)[r1][.]; the zero entry of
      // the skew row comes first in the reference's 3-term sum and does not change the bits
      J[rr][3 + c] = rr == 0 ? __fmaf_rn(cpos[1], Rj(c, 2), __fmul_rn(-cpos[2], Rj(c, 1)))
                   : rr == 1 ? __fmaf_rn(-cpos[0], Rj(c, 2), __fmul_rn(cpos[2], Rj(c, 0)))
                             : __fmaf_rn(cpos[0], Rj(c, 1), __fmul_rn(-cpos[1], Rj(c, 0)));
      J[3 + rr][c] = 0.0f;
      J[3 + rr][3 + c] = rjt;
    }
    J[rr][6] = -rm_tp[rr];
    J[3 + rr][6] = 0.0f;
  }

  // ---- S = J^T W J (upper triangle), v = J^T W r: fp32, residual rows in order ---------------
  double* out = blocks + (size_t)idx * kOdomBlock;
  int l = 0;
#pragma unroll
  for (int a = 0; a < 7; ++a) {
    float jw[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) jw[k] = __fmul_rn(J[k][a], w[k]);
#pragma unroll
    for (int b = a; b < 7; ++b) {
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < 6; ++k) acc = __fmaf_rn(jw[k], J[k][b], acc);
      out[l++] = (double)acc;
    }
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < 6; ++k) acc = __fmaf_rn(jw[k], r[k], acc);
    out[28 + a] = (double)acc;
  }
}

__global__ void __launch_bounds__(64)
odom_add_kernel(const double* __restrict__ blocks, const int32_t* __restrict__ rows, int Eo,
                const float* __restrict__ Twc, const double* __restrict__ s_bar,
                float sigma_scale, int nblk, int num_fix, double* __restrict__ Hg,
                const int32_t* __restrict__ done, const int32_t* __restrict__ pose_rows) {
  if (done != nullptr && *done != 0) return;
  const int bid = blockIdx.x;
  int a = (int)((sqrtf(8.0f * (float)bid + 1.0f) - 1.0f) * 0.5f);
  while ((a + 1) * (a + 2) / 2 <= bid) ++a;
  while (a * (a + 1) / 2 > bid) --a;
  const int b = bid - a * (a + 1) / 2;
  const int n = 7 * nblk;
  const int tid = threadIdx.x;
  const int r = tid / 7, c = tid % 7;
  const bool is_h = tid < 49, is_g = tid >= 49 && tid < 56;
  double acc = 0.0, acc_t = 0.0;  // (7a+r, 7b+c) and its mirror (7b+c, 7a+r)
  // entry (r, c) of S for this thread and of its mirror (c, r): the same number, S is symmetric
  const int lo = min(r, c), hi = max(r, c);
  const int sidx = is_h ? lo * 7 - lo * (lo - 1) / 2 + (hi - lo) : 0;
  const bool pose6 = r < 6 && c < 6;  // entry untouched by the scale row / column
  for (int e = 0; e < Eo; ++e) {
    const int io = rows[e] - num_fix, jo = rows[Eo + e] - num_fix;
    const double* B = blocks + (size_t)e * kOdomBlock;
    if (is_h) {
      const double sv = B[sidx];
      // H_ii = S without its scale row/column; H_ij = -S without its scale ROW (pose i has no
      // scale term, pose j has); H_ji = H_ij^T; H_jj = S
      if (io >= 0 && io == a && io == b && pose6) acc += sv;                         // H_ii
      if (io >= 0 && jo >= 0) {
        if (io == a && jo == b && r < 6) acc -= sv;                                   // H_ij
        if (io == b && jo == a && c < 6) acc_t -= sv;                                 // H_ij, mirror
        if (jo == a && io == b && c < 6) acc -= sv;                                   // H_ji
        if (jo == b && io == a && r < 6) acc_t -= sv;                                 // H_ji, mirror
      }
      if (jo >= 0 && jo == a && jo == b) acc += sv;                                   // H_jj
    } else if (is_g && a == b) {
      const int k = tid - 49;
      if (io == a && k < 6) acc -= B[28 + k];  // g_i = -v without the scale entry
      if (jo == a) acc += B[28 + k];           // g_j = v
    }
  }
  // scale prior on the diagonal block of keyframe row a + num_fix (:1744-1781)
  if (a == b && s_bar != nullptr && (double)sigma_scale > 1e-9 && (tid == 48 || tid == 55)) {
    const int row = a + num_fix;
    const double sbar = s_bar[row];
    const size_t trow = pose_rows ? (size_t)pose_rows[row] : (size_t)row;  // arena: rank -> slot
    const double si = (double)Twc[8 * trow + 6];  // sic: column 6 (q_w), see the header comment
    if (sbar > 1e-6 && si > 1e-6) {
      const double info = 1.0 / ((double)sigma_scale * (double)sigma_scale);
      acc += tid == 48 ? info : info * log(si / sbar);
    }
  }
  if (is_h) {
    Hg[(size_t)(7 * a + r) * n + (7 * b + c)] += acc;
    if (a != b) Hg[(size_t)(7 * b + c) * n + (7 * a + r)] += acc_t;
  } else if (is_g && a == b) {
    Hg[(size_t)n * n + 7 * a + (tid - 49)] += acc;
  }
}

}  // namespace

size_t gn_odom_scratch_bytes(int Eo) {
  const size_t e = (size_t)(Eo > 0 ? Eo : 1);
  const size_t rows = (2 * e * sizeof(int32_t) + 255) / 256 * 256;
  return rows + e * kOdomBlock * sizeof(double);
}

int launch_gn_add_odom(const GnOdom& od, const float* Twc, int K, int num_fix, double* Hg,
                       const int32_t* done, void* scratch, cudaStream_t s) {
  const int32_t* rows = static_cast<const int32_t*>(scratch);
  const size_t rows_bytes =
      (2 * (size_t)(od.Eo > 0 ? od.Eo : 1) * sizeof(int32_t) + 255) / 256 * 256;
  double* blocks = reinterpret_cast<double*>(static_cast<char*>(scratch) + rows_bytes);
  return launch_gn_add_odom_rows(od, Twc, K, num_fix, Hg, done, rows, blocks, nullptr, s);
}

size_t gn_odom_blocks_bytes(int Eo) {
  return (size_t)(Eo > 0 ? Eo : 1) * kOdomBlock * sizeof(double);
}

// rows [2 Eo]: system ranks of the odometry endpoints (ii ranks, then jj ranks); od.odom_ii/jj
// index the rows of Twc; pose_rows (may be NULL): system rank -> row of Twc for the scale prior.
int launch_gn_add_odom_rows(const GnOdom& od, const float* Twc, int K, int num_fix, double* Hg,
                            const int32_t* done, const int32_t* rows, double* blocks,
                            const int32_t* pose_rows, cudaStream_t s) {
  const int nblk = K - num_fix;
  if (nblk <= 0) return FRODO_OK;
  const int Eo = od.Eo > 0 ? od.Eo : 0;
  if (Eo > 0) {
    odom_blocks_kernel<<<(Eo + 63) / 64, 64, 0, s>>>(Twc, od.Tij, od.stride, od.odom_ii,
                                                      od.odom_jj, Eo, od.sigma_t, od.sigma_r,
                                                      blocks, done);
    FRODO_CUDA_TRY(cudaGetLastError());
  }
  const int nb = nblk * (nblk + 1) / 2;
  odom_add_kernel<<<nb, 64, 0, s>>>(blocks, rows, Eo, Twc, od.s_bar, od.sigma_scale, nblk,
                                    num_fix, Hg, done, pose_rows);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
