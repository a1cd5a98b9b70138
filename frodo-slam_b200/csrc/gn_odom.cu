// Extra terms of gauss_newton_rays_odom (gn.cpp:29-62, gn_kernels.cu:1792-1934): relative-pose
// (odometry) constraints between keyframes and a per-keyframe scale prior, added to the dense
// fp64 normal equations after the visual terms have been assembled.
//   * odom_blocks_kernel: one thread per odometry edge, the fp32 arithmetic of
//     odom_constraint_kernel_left_perturb_log (:1106-1286; SE3-log residual of
//     T_meas^-1 * [R_i^T R_j, R_i^T (t_j - t_i) / s_j], anisotropic sigma, Huber, approximate
//     Ad(T_j^-1) Jacobian with a scale column), writing the full 14x14 block + 14-vector;
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

constexpr int kOdomBlock = 14 * 14 + 14;

__device__ __forceinline__ float huber_o(float r) {  // :214-217 (fp64 compare and divide)
  const float r_abs = fabsf(r);
  return (double)r_abs < 1.345 ? 1.0f : (float)(1.345 / (double)r_abs);
}

__device__ void quat_to_rot(const float* q, float* R) {  // :562-574
  const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
  R[0] = 1 - 2 * qy * qy - 2 * qz * qz;
  R[1] = 2 * qx * qy - 2 * qz * qw;
  R[2] = 2 * qx * qz + 2 * qy * qw;
  R[3] = 2 * qx * qy + 2 * qz * qw;
  R[4] = 1 - 2 * qx * qx - 2 * qz * qz;
  R[5] = 2 * qy * qz - 2 * qx * qw;
  R[6] = 2 * qx * qz - 2 * qy * qw;
  R[7] = 2 * qy * qz + 2 * qx * qw;
  R[8] = 1 - 2 * qx * qx - 2 * qy * qy;
}

__device__ void skew(const float* v, float* M) {
  M[0] = 0;     M[1] = -v[2]; M[2] = v[1];
  M[3] = v[2];  M[4] = 0;     M[5] = -v[0];
  M[6] = -v[1]; M[7] = v[0];  M[8] = 0;
}

__device__ void mat_vec(const float* A, const float* x, float* y) {
  y[0] = A[0] * x[0] + A[1] * x[1] + A[2] * x[2];
  y[1] = A[3] * x[0] + A[4] * x[1] + A[5] * x[2];
  y[2] = A[6] * x[0] + A[7] * x[1] + A[8] * x[2];
}

__device__ void mat_mat(const float* A, const float* B, float* C) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      float acc = 0;
      for (int k = 0; k < 3; ++k) acc += A[r * 3 + k] * B[k * 3 + c];
      C[r * 3 + c] = acc;
    }
}

__device__ void transpose(const float* A, float* T) {
  T[0] = A[0]; T[1] = A[3]; T[2] = A[6];
  T[3] = A[1]; T[4] = A[4]; T[5] = A[7];
  T[6] = A[2]; T[7] = A[5]; T[8] = A[8];
}

// SO(3) logarithm, :1007-1041
__device__ void rot_to_vec(const float* R, float* omega) {
  const float trace = R[0] + R[4] + R[8];
  float cos_angle = 0.5f * (trace - 1.0f);
  cos_angle = fmaxf(-1.0f, fminf(1.0f, cos_angle));
  const float angle = acosf(cos_angle);
  if (fabsf(angle) < 1e-6f) {
    omega[0] = omega[1] = omega[2] = 0.0f;
  } else if (fabsf((float)((double)angle - 3.14159265358979323846)) < 1e-6f) {
    if (R[0] > R[4] && R[0] > R[8]) {
      const float sv = 2.0f * sqrtf(R[0] - R[4] - R[8] + 1.0f);
      omega[0] = 0.5f * sv; omega[1] = (R[3] + R[1]) / sv; omega[2] = (R[2] + R[6]) / sv;
    } else if (R[4] > R[8]) {
      const float sv = 2.0f * sqrtf(R[4] - R[0] - R[8] + 1.0f);
      omega[0] = (R[3] + R[1]) / sv; omega[1] = 0.5f * sv; omega[2] = (R[7] + R[5]) / sv;
    } else {
      const float sv = 2.0f * sqrtf(R[8] - R[0] - R[4] + 1.0f);
      omega[0] = (R[2] + R[6]) / sv; omega[1] = (R[7] + R[5]) / sv; omega[2] = 0.5f * sv;
    }
    omega[0] *= angle; omega[1] *= angle; omega[2] *= angle;
  } else {
    const float sin_angle = sinf(angle);
    const float factor = angle / (2.0f * sin_angle);
    omega[0] = factor * (R[7] - R[5]);
    omega[1] = factor * (R[2] - R[6]);
    omega[2] = factor * (R[3] - R[1]);
  }
}

// inverse left Jacobian of SO(3), :1044-1074
__device__ void left_jacobian_inv(const float* omega, float* J_inv) {
  const float angle = sqrtf(omega[0] * omega[0] + omega[1] * omega[1] + omega[2] * omega[2]);
  float sk[9];
  skew(omega, sk);
  for (int i = 0; i < 9; ++i) J_inv[i] = (i % 4 == 0) ? 1.0f : 0.0f;
  if (fabsf(angle) < 1e-6f) {
    for (int i = 0; i < 9; ++i) J_inv[i] -= 0.5f * sk[i];
  } else {
    const float angle_sq = angle * angle;
    const float sin_angle = sinf(angle), cos_angle = cosf(angle);
    const float factor2 = (1.0f / angle_sq) - ((1.0f + cos_angle) / (2.0f * angle * sin_angle));
    float sk2[9];
    mat_mat(sk, sk, sk2);
    for (int i = 0; i < 9; ++i) J_inv[i] += -0.5f * sk[i] + factor2 * sk2[i];
  }
}

__global__ void __launch_bounds__(64)
odom_blocks_kernel(const float* __restrict__ Twc, const float* __restrict__ Tij, int stride,
                   const int64_t* __restrict__ odom_ii, const int64_t* __restrict__ odom_jj,
                   int Eo, float sigma_t, float sigma_r, double* __restrict__ blocks,
                   const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Eo) return;
  const float* Ti = Twc + 8 * odom_ii[idx];
  const float* Tj = Twc + 8 * odom_jj[idx];
  float ti[3], tj[3], qi[4], qj[4], tm[3], qm[4];
  for (int k = 0; k < 3; ++k) { ti[k] = Ti[k]; tj[k] = Tj[k]; tm[k] = Tij[(size_t)stride * idx + k]; }
  for (int k = 0; k < 4; ++k) { qi[k] = Ti[3 + k]; qj[k] = Tj[3 + k]; qm[k] = Tij[(size_t)stride * idx + 3 + k]; }
  const float sj = Tj[7];
  float R_m[9], R_i[9], R_j[9], R_i_T[9], R_j_T[9], R_m_T[9];
  quat_to_rot(qm, R_m);
  quat_to_rot(qi, R_i);
  quat_to_rot(qj, R_j);
  transpose(R_i, R_i_T);
  transpose(R_j, R_j_T);
  transpose(R_m, R_m_T);
  float R_ij[9];
  mat_mat(R_i_T, R_j, R_ij);
  const float t_diff[3] = {tj[0] - ti[0], tj[1] - ti[1], tj[2] - ti[2]};
  float rit[3];
  mat_vec(R_i_T, t_diff, rit);
  const float sj_inv = (sj == 0.0f) ? 1e6f : 1.0f / sj;
  const float t_pred[3] = {sj_inv * rit[0], sj_inv * rit[1], sj_inv * rit[2]};
  float t_meas_inv[3];
  mat_vec(R_m_T, tm, t_meas_inv);
  t_meas_inv[0] = -t_meas_inv[0]; t_meas_inv[1] = -t_meas_inv[1]; t_meas_inv[2] = -t_meas_inv[2];
  float R_err[9], rmt[3], t_err[3];
  mat_mat(R_m_T, R_ij, R_err);
  mat_vec(R_m_T, t_pred, rmt);
  for (int k = 0; k < 3; ++k) t_err[k] = rmt[k] + t_meas_inv[k];
  float r[6], J_inv[9];
  rot_to_vec(R_err, r + 3);
  left_jacobian_inv(r + 3, J_inv);
  mat_vec(J_inv, t_err, r);
  float Wd[6];
  const float info_t = 1.0f / (sigma_t * sigma_t), info_r = 1.0f / (sigma_r * sigma_r);
  for (int k = 0; k < 3; ++k) Wd[k] = info_t * huber_o(r[k] / sigma_t);
  for (int k = 3; k < 6; ++k) Wd[k] = info_r * huber_o(r[k] / sigma_r);
  // J = [J_i | J_j], 6 x 14, J_i = -J_j except for the scale column (zero for pose i)
  float J[84];
  for (int k = 0; k < 84; ++k) J[k] = 0.0f;
  float tj_inv[3], tj_sk[9], drho[9], drho_ds[3];
  mat_vec(R_j_T, tj, tj_inv);
  tj_inv[0] = -tj_inv[0]; tj_inv[1] = -tj_inv[1]; tj_inv[2] = -tj_inv[2];
  skew(tj_inv, tj_sk);
  mat_mat(tj_sk, R_j_T, drho);
  mat_vec(R_m_T, t_pred, drho_ds);
  for (int rr = 0; rr < 3; ++rr) {
    for (int c = 0; c < 3; ++c) {
      J[rr * 14 + 7 + c] = R_j_T[rr * 3 + c];
      J[rr * 14 + 7 + 3 + c] = drho[rr * 3 + c];
      J[(rr + 3) * 14 + 7 + 3 + c] = R_j_T[rr * 3 + c];
    }
    J[rr * 14 + 7 + 6] = -drho_ds[rr];
  }
  for (int rr = 0; rr < 6; ++rr)
    for (int c = 0; c < 6; ++c) J[rr * 14 + c] = -J[rr * 14 + 7 + c];
  double* out = blocks + (size_t)idx * kOdomBlock;
  for (int a = 0; a < 14; ++a) {
    float jtw[6];
    for (int k = 0; k < 6; ++k) jtw[k] = J[k * 14 + a] * Wd[k];
    for (int b = 0; b < 14; ++b) {
      float val = 0;
      for (int k = 0; k < 6; ++k) val += jtw[k] * J[k * 14 + b];
      out[a * 14 + b] = (double)val;
    }
    float val = 0;
    for (int k = 0; k < 6; ++k) val += jtw[k] * r[k];
    out[196 + a] = (double)val;
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
  for (int e = 0; e < Eo; ++e) {
    const int io = rows[e] - num_fix, jo = rows[Eo + e] - num_fix;
    const double* B = blocks + (size_t)e * kOdomBlock;
    if (is_h) {
      if (io >= 0) {
        if (io == a && io == b) acc += B[r * 14 + c];                       // H_ii
        if (jo >= 0 && io == a && jo == b) acc += B[r * 14 + 7 + c];        // H_ij
        if (jo >= 0 && io == b && jo == a) acc_t += B[c * 14 + 7 + r];      // H_ij, mirror
      }
      if (jo >= 0) {
        if (io >= 0 && jo == a && io == b) acc += B[(7 + r) * 14 + c];      // H_ji
        if (io >= 0 && jo == b && io == a) acc_t += B[(7 + c) * 14 + r];    // H_ji, mirror
        if (jo == a && jo == b) acc += B[(7 + r) * 14 + 7 + c];             // H_jj
      }
    } else if (is_g && a == b) {
      if (io == a) acc += B[196 + (tid - 49)];
      if (jo == a) acc += B[196 + 7 + (tid - 49)];
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
