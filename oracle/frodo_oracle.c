/*
 * frodo_oracle.c -- CPU restatement of the mast3r_slam_backends hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (frodo-slam_b200/) may
 * import, link or execute this file; it is the parity checker used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * Parity status: the reference ships no tests or golden vectors for this path
 * (SURVEY.md section 4), so the oracle is pinned against outputs of the
 * reference's own kernels (oracle/_ref, built by oracle/build_ref.sh) recorded
 * on a B200 by tests/golden/make_golden_ref.py and committed under
 * tests/golden/.  See DESIGN.md "Oracle".
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/mast3r_slam/backend/src/).  The restatement is scalar and
 * follows the reference expression by expression, including
 *   - the implicit fp64 sub-expressions created by unsuffixed literals,
 *   - the FMA contractions nvcc applies to the reference source (verified
 *     against the SASS of oracle/_ref; written here as explicit fmaf()),
 *   - fp16 round-to-nearest-even after every multiply and every add in
 *     refine_matches (c10::Half arithmetic).
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fno-fast-math (see oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

ORACLE_API int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

ORACLE_API void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ------------------------------------------------------------------------ */
/* fp16 <-> fp32 (IEEE binary16, round-to-nearest-even, subnormals, inf/nan) */
/* ------------------------------------------------------------------------ */

static inline float h2f(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu;
  uint32_t man = h & 0x3FFu;
  uint32_t bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else { /* subnormal: value = man * 2^-24 */
      float f = (float)man * 5.9604644775390625e-08f;
      memcpy(&bits, &f, 4);
      bits |= sign;
    }
  } else if (exp == 31) {
    bits = sign | 0x7F800000u | (man << 13);
  } else {
    bits = sign | ((exp + 112u) << 23) | (man << 13);
  }
  float out;
  memcpy(&out, &bits, 4);
  return out;
}

static inline uint16_t f2h(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  uint32_t sign = (x >> 16) & 0x8000u;
  uint32_t ax = x & 0x7FFFFFFFu;
  if (ax >= 0x7F800000u) { /* inf / nan */
    return (uint16_t)(sign | 0x7C00u | ((ax > 0x7F800000u) ? 0x200u : 0u));
  }
  if (ax >= 0x477FF000u) { /* >= 65520 rounds to inf */
    return (uint16_t)(sign | 0x7C00u);
  }
  if (ax < 0x38800000u) { /* < 2^-14: subnormal half (or zero) */
    if (ax < 0x33000000u) { /* < 2^-25 : rounds to zero (2^-25 itself ties to even = 0) */
      return (uint16_t)sign;
    }
    /* value = m * 2^(e-150); want round(value / 2^-24) */
    uint32_t e = ax >> 23;
    uint32_t m = (ax & 0x7FFFFFu) | 0x800000u;
    uint32_t shift = 126u - e; /* in [14, 24] */
    uint32_t q = m >> shift;
    uint32_t rem = m & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1u);
    if (rem > half || (rem == half && (q & 1u))) q++;
    return (uint16_t)(sign | q);
  }
  /* normal */
  uint32_t e = (ax >> 23) - 112u;
  uint32_t m = ax & 0x7FFFFFu;
  uint32_t q = (e << 10) | (m >> 13);
  uint32_t rem = m & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
  return (uint16_t)(sign | q);
}

ORACLE_API void oracle_f2h(const float* in, uint16_t* out, int64_t n) {
  for (int64_t i = 0; i < n; i++) out[i] = f2h(in[i]);
}
ORACLE_API void oracle_h2f(const uint16_t* in, float* out, int64_t n) {
  for (int64_t i = 0; i < n; i++) out[i] = h2f(in[i]);
}

/* ------------------------------------------------------------------------ */
/* iter_proj   (matching_kernels.cu:128-284)                                 */
/* ------------------------------------------------------------------------ */

/* matching_kernels.cu:21-23 : fmin(fmax(x, lo), hi) on floats */
static inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

/* FMA contraction as found in the SASS of the reference build (oracle/_ref,
 * nvcc 12.9 -O3 sm_100; cuobjdump -sass of iter_proj_kernel), see DESIGN.md:
 *   a0*b0 + a1*b1 + a2*b2          -> fma(a2,b2, fma(a0,b0, a1*b1))
 *   w11*a + w12*b + w21*c + w22*d  -> fma(w22,d, fma(w21,c, fma(w11,a, w12*b)))
 *   r[j]*=inv; err[j]=r[j]-t[j]    -> err[j] = fma(r[j], inv, -t[j])
 *   A00*A11 - A01*A01              -> fma(A00, A11, -(A01*A01))
 *   u + det_inv*(A11*b0 - A01*b1)  -> fma(fma(A01, e.gy, -(e.gx*A11)), det_inv, u)
 *   v + det_inv*(-A01*b0 + A00*b1) -> fma(fma(A01, e.gx, -(e.gy*A00)), det_inv, v)
 *   1.0/x (fp64) -> float          == IEEE fp32 1/x  (the compiler emits exactly that)
 */
static inline float blend4(float w11, float a, float w12, float b, float w21, float c, float w22,
                           float d) {
  return fmaf(w22, d, fmaf(w21, c, fmaf(w11, a, w12 * b)));
}

static inline float dot3c(float ax, float ay, float az, float bx, float by, float bz) {
  return fmaf(az, bz, fmaf(ax, bx, ay * by));
}

ORACLE_API void oracle_iter_proj(const float* rays, /* [B,H,W,9] */
                                 const float* pts,  /* [B,N,3]   */
                                 const float* p_init, /* [B,N,2] */
                                 float* p_new,        /* [B,N,2] */
                                 uint8_t* converged,  /* [B,N]   */
                                 int B, int H, int W, int N, int max_iter, float lambda_init,
                                 float cost_thresh) {
  const int64_t total = (int64_t)B * N;
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < total; t++) {
    const int64_t b = t / N;
    const float* img = rays + (size_t)b * H * W * 9;
    const float tx = pts[t * 3 + 0], ty = pts[t * 3 + 1], tz = pts[t * 3 + 2];
    /* :148-153 */
    float u = clampf(p_init[t * 2 + 0], 1.0f, (float)(W - 2));
    float v = clampf(p_init[t * 2 + 1], 1.0f, (float)(H - 2));
    float lambda = lambda_init;
    uint8_t conv = 0; /* torch::zeros, matching_kernels.cu:309 */
    for (int it = 0; it < max_iter; it++) {
      /* :164-173 -- (1.0-du) etc. are fp64 in the reference; du,dv are multiples
       * of 2^-23 in [0,1) so 1-du is exact and the fp64 products round once. */
      int u11 = (int)floorf(u);
      int v11 = (int)floorf(v);
      float du = u - (float)u11;
      float dv = v - (float)v11;
      float w11 = du * dv;
      float w12 = (float)((1.0 - (double)du) * (double)dv);
      float w21 = (float)((double)du * (1.0 - (double)dv));
      float w22 = (float)((1.0 - (double)du) * (1.0 - (double)dv));
      /* :176-179  (names are crossed on purpose in the reference) */
      const float* r11 = img + ((size_t)(v11 + 1) * W + (u11 + 1)) * 9;
      const float* r12 = img + ((size_t)(v11 + 1) * W + u11) * 9;
      const float* r21 = img + ((size_t)v11 * W + (u11 + 1)) * 9;
      const float* r22 = img + ((size_t)v11 * W + u11) * 9;
      float r[3], gx[3], gy[3], err[3];
      for (int j = 0; j < 3; j++) r[j] = blend4(w11, r11[j], w12, r12[j], w21, r21[j], w22, r22[j]);
      for (int j = 0; j < 3; j++)
        gx[j] = blend4(w11, r11[3 + j], w12, r12[3 + j], w21, r21[3 + j], w22, r22[3 + j]);
      for (int j = 0; j < 3; j++)
        gy[j] = blend4(w11, r11[6 + j], w12, r12[6 + j], w21, r21[6 + j], w22, r22[6 + j]);
      /* :195-200 */
      float r_norm = sqrtf(dot3c(r[0], r[1], r[2], r[0], r[1], r[2]));
      float r_norm_inv = (float)(1.0 / (double)r_norm);
      /* :198-206 (scale and subtract are fused by nvcc) */
      err[0] = fmaf(r[0], r_norm_inv, -tx);
      err[1] = fmaf(r[1], r_norm_inv, -ty);
      err[2] = fmaf(r[2], r_norm_inv, -tz);
      float cost = dot3c(err[0], err[1], err[2], err[0], err[1], err[2]);
      /* :211-219 */
      float A00 = dot3c(gx[0], gx[1], gx[2], gx[0], gx[1], gx[2]);
      float A01 = dot3c(gx[0], gx[1], gx[2], gy[0], gy[1], gy[2]);
      float A11 = dot3c(gy[0], gy[1], gy[2], gy[0], gy[1], gy[2]);
      float egx = dot3c(err[0], err[1], err[2], gx[0], gx[1], gx[2]); /* = -b0 */
      float egy = dot3c(err[0], err[1], err[2], gy[0], gy[1], gy[2]); /* = -b1 */
      A00 += lambda;
      A11 += lambda;
      /* :222-230 */
      float det = fmaf(A00, A11, -(A01 * A01));
      float det_inv = (float)(1.0 / (double)det);
      float num_u = fmaf(A01, egy, -(egx * A11)); /* A11*b0 - A01*b1 */
      float num_v = fmaf(A01, egx, -(egy * A00)); /* -A01*b0 + A00*b1 */
      float u_new = clampf(fmaf(num_u, det_inv, u), 1.0f, (float)(W - 2));
      float v_new = clampf(fmaf(num_v, det_inv, v), 1.0f, (float)(H - 2));
      /* :234-265 */
      u11 = (int)floorf(u_new);
      v11 = (int)floorf(v_new);
      du = u_new - (float)u11;
      dv = v_new - (float)v11;
      w11 = du * dv;
      w12 = (float)((1.0 - (double)du) * (double)dv);
      w21 = (float)((double)du * (1.0 - (double)dv));
      w22 = (float)((1.0 - (double)du) * (1.0 - (double)dv));
      r11 = img + ((size_t)(v11 + 1) * W + (u11 + 1)) * 9;
      r12 = img + ((size_t)(v11 + 1) * W + u11) * 9;
      r21 = img + ((size_t)v11 * W + (u11 + 1)) * 9;
      r22 = img + ((size_t)v11 * W + u11) * 9;
      for (int j = 0; j < 3; j++) r[j] = blend4(w11, r11[j], w12, r12[j], w21, r21[j], w22, r22[j]);
      r_norm = sqrtf(dot3c(r[0], r[1], r[2], r[0], r[1], r[2]));
      r_norm_inv = (float)(1.0 / (double)r_norm);
      err[0] = fmaf(r[0], r_norm_inv, -tx);
      err[1] = fmaf(r[1], r_norm_inv, -ty);
      err[2] = fmaf(r[2], r_norm_inv, -tz);
      float new_cost = dot3c(err[0], err[1], err[2], err[0], err[1], err[2]);
      /* :268-277 */
      if (new_cost < cost) {
        u = u_new;
        v = v_new;
        lambda = (float)((double)lambda * 0.1);
        conv = new_cost < cost_thresh;
      } else {
        lambda = (float)((double)lambda * 10.0);
        conv = cost < cost_thresh;
      }
    }
    p_new[t * 2 + 0] = u;
    p_new[t * 2 + 1] = v;
    converged[t] = conv;
  }
}

/* ------------------------------------------------------------------------ */
/* refine_matches   (matching_kernels.cu:25-85)                              */
/* ------------------------------------------------------------------------ */

/* scalar_t = c10::Half: every '*' and '+=' is float math rounded to fp16
 * (torch/headeronly/util/Half.h); max_score starts at
 * cuda::std::numeric_limits<c10::Half>::min() == Half() == 0 (:48). */
ORACLE_API void oracle_refine_matches_f16(const uint16_t* D11, /* [B,H,W,F] */
                                          const uint16_t* D21, /* [B,N,F]   */
                                          const int64_t* p1,   /* [B,N,2]   */
                                          int64_t* p1_new,     /* [B,N,2]   */
                                          uint16_t* max_scores, /* [B,N]    */
                                          int B, int H, int W, int N, int F, int radius,
                                          int dilation_max) {
  const int64_t total = (int64_t)B * N;
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < total; t++) {
    const int64_t b = t / N;
    const uint16_t* img = D11 + (size_t)b * H * W * F;
    const uint16_t* q = D21 + (size_t)t * F;
    int64_t u0 = p1[t * 2 + 0], v0 = p1[t * 2 + 1];
    uint16_t max_score = 0;
    float max_score_f = 0.0f;
    int64_t u_new = u0, v_new = v0;
    for (int d = dilation_max; d > 0; d--) {
      const int rd = radius * d;
      const int diam = 2 * rd + 1;
      for (int i = 0; i < diam; i += d) {
        for (int j = 0; j < diam; j += d) {
          const int64_t u = u0 - rd + i;
          const int64_t v = v0 - rd + j;
          if (v >= 0 && v < H && u >= 0 && u < W) { /* :17-19 (int-truncated in ref) */
            const uint16_t* c = img + ((size_t)v * W + (size_t)u) * F;
            uint16_t score = 0;
            for (int k = 0; k < F; k++) {
              uint16_t prod = f2h(h2f(q[k]) * h2f(c[k]));
              score = f2h(h2f(score) + h2f(prod));
            }
            float sf = h2f(score);
            if (sf > max_score_f) {
              max_score = score;
              max_score_f = sf;
              u_new = u;
              v_new = v;
            }
          }
        }
      }
      u0 = u_new;
      v0 = v_new;
    }
    p1_new[t * 2 + 0] = u_new;
    p1_new[t * 2 + 1] = v_new;
    max_scores[t] = max_score;
  }
}

/* scalar_t = float: plain fp32 with nvcc's FMA contraction of
 * score += a*b  ->  score = fma(a, b, score); max_score starts at FLT_MIN. */
ORACLE_API void oracle_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                                          int64_t* p1_new, float* max_scores, int B, int H, int W,
                                          int N, int F, int radius, int dilation_max) {
  const int64_t total = (int64_t)B * N;
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < total; t++) {
    const int64_t b = t / N;
    const float* img = D11 + (size_t)b * H * W * F;
    const float* q = D21 + (size_t)t * F;
    int64_t u0 = p1[t * 2 + 0], v0 = p1[t * 2 + 1];
    float max_score = 1.17549435e-38f;
    int64_t u_new = u0, v_new = v0;
    for (int d = dilation_max; d > 0; d--) {
      const int rd = radius * d;
      const int diam = 2 * rd + 1;
      for (int i = 0; i < diam; i += d) {
        for (int j = 0; j < diam; j += d) {
          const int64_t u = u0 - rd + i;
          const int64_t v = v0 - rd + j;
          if (v >= 0 && v < H && u >= 0 && u < W) {
            const float* c = img + ((size_t)v * W + (size_t)u) * F;
            float score = 0.0f;
            for (int k = 0; k < F; k++) score = fmaf(q[k], c[k], score);
            if (score > max_score) {
              max_score = score;
              u_new = u;
              v_new = v;
            }
          }
        }
      }
      u0 = u_new;
      v0 = v_new;
    }
    p1_new[t * 2 + 0] = u_new;
    p1_new[t * 2 + 1] = v_new;
    max_scores[t] = max_score;
  }
}

/* ------------------------------------------------------------------------ */
/* Sim3 helpers   (gn_kernels.cu:214-455)                                    */
/* ------------------------------------------------------------------------ */

/* :214-217  fabs(r) < 1.345 (fp64 compare) ? 1.0 : 1.345/r_abs (fp64 divide) */
static inline float huber(float r) {
  const float r_abs = fabsf(r);
  return ((double)r_abs < 1.345) ? 1.0f : (float)(1.345 / (double)r_abs);
}

/* :220-226 */
static void quat_comp(const float* qi, const float* qj, float* out) {
  out[0] = qi[3] * qj[0] + qi[0] * qj[3] + qi[1] * qj[2] - qi[2] * qj[1];
  out[1] = qi[3] * qj[1] - qi[0] * qj[2] + qi[1] * qj[3] + qi[2] * qj[0];
  out[2] = qi[3] * qj[2] + qi[0] * qj[1] - qi[1] * qj[0] + qi[2] * qj[3];
  out[3] = qi[3] * qj[3] - qi[0] * qj[0] - qi[1] * qj[1] - qi[2] * qj[2];
}

/* :237-247 (2.0* promotes to fp64, result stored to float) */
static void actSO3(const float* q, const float* X, float* Y) {
  float uv[3];
  uv[0] = (float)(2.0 * (double)(q[1] * X[2] - q[2] * X[1]));
  uv[1] = (float)(2.0 * (double)(q[2] * X[0] - q[0] * X[2]));
  uv[2] = (float)(2.0 * (double)(q[0] * X[1] - q[1] * X[0]));
  float y0 = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  float y1 = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  float y2 = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
  Y[0] = y0;
  Y[1] = y1;
  Y[2] = y2;
}

/* :249-261 */
static void actSim3(const float* t, const float* q, const float* s, const float* X, float* Y) {
  actSO3(q, X, Y);
  Y[0] *= s[0];
  Y[1] *= s[0];
  Y[2] *= s[0];
  Y[0] += t[0];
  Y[1] += t[1];
  Y[2] += t[2];
}

/* :294-314 */
static void relSim3(const float* ti, const float* qi, const float* si, const float* tj,
                    const float* qj, const float* sj, float* tij, float* qij, float* sij) {
  float si_inv = (float)(1.0 / (double)si[0]);
  sij[0] = si_inv * sj[0];
  float qi_inv[4] = {-qi[0], -qi[1], -qi[2], qi[3]};
  quat_comp(qi_inv, qj, qij);
  float d[3] = {tj[0] - ti[0], tj[1] - ti[1], tj[2] - ti[2]};
  actSO3(qi_inv, d, tij);
  tij[0] *= si_inv;
  tij[1] *= si_inv;
  tij[2] *= si_inv;
}

/* :319-339 */
static void apply_Sim3_adj_inv(const float* t, const float* q, const float* s, const float* X,
                               float* Y) {
  const float s_inv = (float)(1.0 / (double)s[0]);
  float Ra[3];
  actSO3(q, &X[0], Ra);
  Y[0] = s_inv * Ra[0];
  Y[1] = s_inv * Ra[1];
  Y[2] = s_inv * Ra[2];
  actSO3(q, &X[3], &Y[3]);
  Y[3] += s_inv * (t[1] * Ra[2] - t[2] * Ra[1]);
  Y[4] += s_inv * (t[2] * Ra[0] - t[0] * Ra[2]);
  Y[5] += s_inv * (t[0] * Ra[1] - t[1] * Ra[0]);
  Y[6] = X[6] + (s_inv * (t[0] * Ra[0] + t[1] * Ra[1] + t[2] * Ra[2]));
}

#define ORACLE_EPS 1e-6

/* :341-363 */
static void expSO3(const float* phi, float* q) {
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float imag, real;
  if ((double)theta_sq < ORACLE_EPS) {
    float theta_p4 = theta_sq * theta_sq;
    imag = (float)(0.5 - (1.0 / 48.0) * (double)theta_sq + (1.0 / 3840.0) * (double)theta_p4);
    real = (float)(1.0 - (1.0 / 8.0) * (double)theta_sq + (1.0 / 384.0) * (double)theta_p4);
  } else {
    float theta = sqrtf(theta_sq);
    imag = sinf((float)(0.5 * (double)theta)) / theta;
    real = cosf((float)(0.5 * (double)theta));
  }
  q[0] = imag * phi[0];
  q[1] = imag * phi[1];
  q[2] = imag * phi[2];
  q[3] = real;
}

static void cross_inplace(const float* a, float* b) { /* :271-282 */
  float x0 = a[1] * b[2] - a[2] * b[1];
  float x1 = a[2] * b[0] - a[0] * b[2];
  float x2 = a[0] * b[1] - a[1] * b[0];
  b[0] = x0;
  b[1] = x1;
  b[2] = x2;
}

/* :365-432 */
static void expSim3(const float* xi, float* t, float* q, float* s) {
  float tau[3] = {xi[0], xi[1], xi[2]};
  float phi[3] = {xi[3], xi[4], xi[5]};
  float sigma = xi[6];
  float scale = expf(sigma);
  expSO3(phi, q);
  s[0] = scale;
  float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  float theta = sqrtf(theta_sq);
  float A, Bc, C;
  const float one = 1.0f, half = 0.5f;
  if ((double)fabsf(sigma) < ORACLE_EPS) {
    C = one;
    if ((double)fabsf(theta) < ORACLE_EPS) {
      A = half;
      Bc = (float)(1.0 / 6.0);
    } else {
      A = (one - cosf(theta)) / theta_sq;
      Bc = (theta - sinf(theta)) / (theta_sq * theta);
    }
  } else {
    C = (scale - one) / sigma;
    if ((double)fabsf(theta) < ORACLE_EPS) {
      float sigma_sq = sigma * sigma;
      A = ((sigma - one) * scale + one) / sigma_sq;
      Bc = (scale * half * sigma_sq + scale - one - sigma * scale) / (sigma_sq * sigma);
    } else {
      float a = scale * sinf(theta);
      float b = scale * cosf(theta);
      float c = theta_sq + sigma * sigma;
      A = (a * sigma + (one - b) * theta) / (theta * c);
      Bc = (C - ((b - one) * sigma + a * theta) / (c)) / (theta_sq);
    }
  }
  t[0] = C * tau[0];
  t[1] = C * tau[1];
  t[2] = C * tau[2];
  cross_inplace(phi, tau);
  t[0] += A * tau[0];
  t[1] += A * tau[1];
  t[2] += A * tau[2];
  cross_inplace(phi, tau);
  t[0] += Bc * tau[0];
  t[1] += Bc * tau[1];
  t[2] += Bc * tau[2];
}

/* :434-455 + pose_retr_kernel :522-560 */
ORACLE_API void oracle_pose_retr(float* Twc /*[K,8]*/, const float* dx /*[K-num_fix,7]*/, int K,
                                 int num_fix) {
  for (int k = num_fix; k < K; k++) {
    float* P = Twc + (size_t)k * 8;
    float t[3] = {P[0], P[1], P[2]};
    float q[4] = {P[3], P[4], P[5], P[6]};
    float s = P[7];
    float dt[3] = {0, 0, 0}, dq[4] = {0, 0, 0, 1}, ds[1] = {0};
    expSim3(dx + (size_t)(k - num_fix) * 7, dt, dq, ds);
    float q1[4], t1[3];
    quat_comp(dq, q, q1);
    actSO3(dq, t, t1);
    t1[0] *= ds[0];
    t1[1] *= ds[0];
    t1[2] *= ds[0];
    t1[0] += dt[0];
    t1[1] += dt[1];
    t1[2] += dt[2];
    P[0] = t1[0];
    P[1] = t1[1];
    P[2] = t1[2];
    P[3] = q1[0];
    P[4] = q1[1];
    P[5] = q1[2];
    P[6] = q1[3];
    P[7] = ds[0] * s;
  }
}

/* ------------------------------------------------------------------------ */
/* per-edge residual / Jacobian / Hessian blocks                             */
/*   ray_align_kernel   gn_kernels.cu:576-907                                */
/*   point_align_kernel gn_kernels.cu:1936-2204                              */
/*   calib_proj_kernel  gn_kernels.cu:2295-2607                              */
/* Element math in fp32 as in the reference; the sums over points are kept  */
/* in fp64 (the reference sums in fp32 with a thread-strided order; parity  */
/* is norm-wise 1e-4).                                            */
/* ------------------------------------------------------------------------ */

typedef struct {
  int mode; /* 0 rays, 1 points, 2 calib */
  float sigma_a; /* rays: sigma_ray;  points: sigma_point; calib: sigma_pixel */
  float sigma_b; /* rays: sigma_dist;                      calib: sigma_depth */
  float C_thresh, Q_thresh;
  /* calib only */
  float fx, fy, cx, cy;
  int height, width, pixel_border;
  float z_eps;
} gn_params_t;

/* accumulate one residual row: hij(105 upper-tri of 14x14, row-major n>=m),
 * vi, vj  -- gn_kernels.cu:768-782 */
static inline void accum_row(const float* ti, const float* qi, const float* si, float* Jx, float w,
                             float err, double* hij, double* vi, double* vj) {
  float* Ji = &Jx[0];
  float* Jj = &Jx[7];
  apply_Sim3_adj_inv(ti, qi, si, Ji, Jj);
  for (int n = 0; n < 7; n++) Ji[n] = -Jj[n];
  int l = 0;
  for (int n = 0; n < 14; n++) {
    for (int m = 0; m <= n; m++) {
      hij[l] += (double)(w * Jx[n] * Jx[m]);
      l++;
    }
  }
  for (int n = 0; n < 7; n++) {
    vi[n] += (double)(w * err * Ji[n]);
    vj[n] += (double)(w * err * Jj[n]);
  }
}

static void edge_blocks(const gn_params_t* P, const float* Twc, const float* Xs, const float* Cs,
                        int ix, int jx, const int64_t* idx, const uint8_t* valid_match,
                        const float* Q, int N, double* Hs4 /*[4][49]*/, double* gs2 /*[2][7]*/) {
  const float* Ti = Twc + (size_t)ix * 8;
  const float* Tj = Twc + (size_t)jx * 8;
  const float* ti = Ti;
  const float* qi = Ti + 3;
  const float* si = Ti + 7;
  float tij[3], qij[4], sij[1];
  relSim3(ti, qi, si, Tj, Tj + 3, Tj + 7, tij, qij, sij);

  double hij[105];
  double vi[7], vj[7];
  memset(hij, 0, sizeof hij);
  memset(vi, 0, sizeof vi);
  memset(vj, 0, sizeof vj);

  const float sa_inv = (float)(1.0 / (double)P->sigma_a);
  const float sb_inv = (P->mode == 1) ? 0.0f : (float)(1.0 / (double)P->sigma_b);
  const float* Xi_base = Xs + (size_t)ix * N * 3;
  const float* Xj_base = Xs + (size_t)jx * N * 3;
  const float* Ci_base = Cs + (size_t)ix * N;
  const float* Cj_base = Cs + (size_t)jx * N;

  for (int k = 0; k < N; k++) {
    const int vm = valid_match[k] != 0;
    const int64_t ind = vm ? idx[k] : 0;
    const float Xi[3] = {Xi_base[ind * 3], Xi_base[ind * 3 + 1], Xi_base[ind * 3 + 2]};
    const float Xj[3] = {Xj_base[(size_t)k * 3], Xj_base[(size_t)k * 3 + 1],
                         Xj_base[(size_t)k * 3 + 2]};
    float Pj[3];
    actSim3(tij, qij, sij, Xj, Pj);
    const float q = Q[k];
    const float ci = Ci_base[ind];
    const float cj = Cj_base[k];
    int valid = vm & (q > P->Q_thresh) & (ci > P->C_thresh) & (cj > P->C_thresh);
    float Jx[14];
    float* Ji = Jx;

    if (P->mode == 0) { /* rays: gn_kernels.cu:687-858 */
      const float norm2_i = Xi[0] * Xi[0] + Xi[1] * Xi[1] + Xi[2] * Xi[2];
      const float norm1_i = sqrtf(norm2_i);
      const float norm1_i_inv = (float)(1.0 / (double)norm1_i);
      float ri[3] = {norm1_i_inv * Xi[0], norm1_i_inv * Xi[1], norm1_i_inv * Xi[2]};
      const float norm2_j = Pj[0] * Pj[0] + Pj[1] * Pj[1] + Pj[2] * Pj[2];
      const float norm1_j = sqrtf(norm2_j);
      const float norm1_j_inv = (float)(1.0 / (double)norm1_j);
      float rj[3] = {norm1_j_inv * Pj[0], norm1_j_inv * Pj[1], norm1_j_inv * Pj[2]};
      float err[4] = {rj[0] - ri[0], rj[1] - ri[1], rj[2] - ri[2], norm1_j - norm1_i};
      const float sw_ray = valid ? sa_inv * sqrtf(q) : 0.0f;
      const float sw_dist = valid ? sb_inv * sqrtf(q) : 0.0f;
      float w[4];
      w[0] = huber(sw_ray * err[0]) * (sw_ray * sw_ray);
      w[1] = huber(sw_ray * err[1]) * (sw_ray * sw_ray);
      w[2] = huber(sw_ray * err[2]) * (sw_ray * sw_ray);
      w[3] = huber(sw_dist * err[3]) * (sw_dist * sw_dist);
      const float norm3_j_inv = norm1_j_inv / norm2_j;
      const float drx_dPx = norm1_j_inv - Pj[0] * Pj[0] * norm3_j_inv;
      const float dry_dPy = norm1_j_inv - Pj[1] * Pj[1] * norm3_j_inv;
      const float drz_dPz = norm1_j_inv - Pj[2] * Pj[2] * norm3_j_inv;
      const float drx_dPy = -Pj[0] * Pj[1] * norm3_j_inv;
      const float drx_dPz = -Pj[0] * Pj[2] * norm3_j_inv;
      const float dry_dPz = -Pj[1] * Pj[2] * norm3_j_inv;
      Ji[0] = drx_dPx; Ji[1] = drx_dPy; Ji[2] = drx_dPz; Ji[3] = 0.0f; Ji[4] = rj[2]; Ji[5] = -rj[1]; Ji[6] = 0.0f;
      accum_row(ti, qi, si, Jx, w[0], err[0], hij, vi, vj);
      Ji[0] = drx_dPy; Ji[1] = dry_dPy; Ji[2] = dry_dPz; Ji[3] = -rj[2]; Ji[4] = 0.0f; Ji[5] = rj[0]; Ji[6] = 0.0f;
      accum_row(ti, qi, si, Jx, w[1], err[1], hij, vi, vj);
      Ji[0] = drx_dPz; Ji[1] = dry_dPz; Ji[2] = drz_dPz; Ji[3] = rj[1]; Ji[4] = -rj[0]; Ji[5] = 0.0f; Ji[6] = 0.0f;
      accum_row(ti, qi, si, Jx, w[2], err[2], hij, vi, vj);
      Ji[0] = rj[0]; Ji[1] = rj[1]; Ji[2] = rj[2]; Ji[3] = 0.0f; Ji[4] = 0.0f; Ji[5] = 0.0f; Ji[6] = norm1_j;
      accum_row(ti, qi, si, Jx, w[3], err[3], hij, vi, vj);
    } else if (P->mode == 1) { /* points: gn_kernels.cu:2045-2155 */
      float err[3] = {Pj[0] - Xi[0], Pj[1] - Xi[1], Pj[2] - Xi[2]};
      const float sw = valid ? sa_inv * sqrtf(q) : 0.0f;
      float w[3];
      w[0] = huber(sw * err[0]) * (sw * sw);
      w[1] = huber(sw * err[1]) * (sw * sw);
      w[2] = huber(sw * err[2]) * (sw * sw);
      Ji[0] = 1.0f; Ji[1] = 0.0f; Ji[2] = 0.0f; Ji[3] = 0.0f; Ji[4] = Pj[2]; Ji[5] = -Pj[1]; Ji[6] = Pj[0];
      accum_row(ti, qi, si, Jx, w[0], err[0], hij, vi, vj);
      Ji[0] = 0.0f; Ji[1] = 1.0f; Ji[2] = 0.0f; Ji[3] = -Pj[2]; Ji[4] = 0.0f; Ji[5] = Pj[0]; Ji[6] = Pj[1];
      accum_row(ti, qi, si, Jx, w[1], err[1], hij, vi, vj);
      Ji[0] = 0.0f; Ji[1] = 0.0f; Ji[2] = 1.0f; Ji[3] = Pj[1]; Ji[4] = -Pj[0]; Ji[5] = 0.0f; Ji[6] = Pj[2];
      accum_row(ti, qi, si, Jx, w[2], err[2], hij, vi, vj);
    } else { /* calib: gn_kernels.cu:2424-2559 */
      const int u_target = (int)(ind % P->width);
      const int v_target = (int)(ind / P->width);
      const int valid_z = (Pj[2] > P->z_eps) && (Xi[2] > P->z_eps);
      const float zj_inv = valid_z ? (float)(1.0 / (double)Pj[2]) : 0.0f;
      const float zj_log = valid_z ? logf(Pj[2]) : 0.0f;
      const float zi_log = valid_z ? logf(Xi[2]) : 0.0f;
      const float x_div_z = Pj[0] * zj_inv;
      const float y_div_z = Pj[1] * zj_inv;
      const float u = P->fx * x_div_z + P->cx;
      const float v = P->fy * y_div_z + P->cy;
      const int valid_u = (u > (float)P->pixel_border) && (u < (float)(P->width - 1 - P->pixel_border));
      const int valid_v = (v > (float)P->pixel_border) && (v < (float)(P->height - 1 - P->pixel_border));
      float err[3] = {u - (float)u_target, v - (float)v_target, zj_log - zi_log};
      valid = valid & valid_u & valid_v & valid_z;
      const float sw_px = valid ? sa_inv * sqrtf(q) : 0.0f;
      const float sw_d = valid ? sb_inv * sqrtf(q) : 0.0f;
      float w[3];
      w[0] = huber(sw_px * err[0]) * (sw_px * sw_px);
      w[1] = huber(sw_px * err[1]) * (sw_px * sw_px);
      w[2] = huber(sw_d * err[2]) * (sw_d * sw_d);
      const float fx = P->fx, fy = P->fy;
      Ji[0] = fx * zj_inv; Ji[1] = 0.0f; Ji[2] = -fx * x_div_z * zj_inv; Ji[3] = -fx * x_div_z * y_div_z;
      Ji[4] = fx * (1 + x_div_z * x_div_z); Ji[5] = -fx * y_div_z; Ji[6] = 0.0f;
      accum_row(ti, qi, si, Jx, w[0], err[0], hij, vi, vj);
      Ji[0] = 0.0f; Ji[1] = fy * zj_inv; Ji[2] = -fy * y_div_z * zj_inv; Ji[3] = -fy * (1 + y_div_z * y_div_z);
      Ji[4] = fy * x_div_z * y_div_z; Ji[5] = fy * x_div_z; Ji[6] = 0.0f;
      accum_row(ti, qi, si, Jx, w[1], err[1], hij, vi, vj);
      Ji[0] = 0.0f; Ji[1] = 0.0f; Ji[2] = zj_inv; Ji[3] = y_div_z; Ji[4] = -x_div_z; Ji[5] = 0.0f; Ji[6] = 1.0f;
      accum_row(ti, qi, si, Jx, w[2], err[2], hij, vi, vj);
    }
  }

  /* fill-in: gn_kernels.cu:883-906 */
  for (int n = 0; n < 7; n++) {
    gs2[n] = vi[n];
    gs2[7 + n] = vj[n];
  }
  int l = 0;
  for (int n = 0; n < 14; n++) {
    for (int m = 0; m <= n; m++) {
      double val = hij[l];
      if (n < 7 && m < 7) {
        Hs4[0 * 49 + n * 7 + m] = val;
        Hs4[0 * 49 + m * 7 + n] = val;
      } else if (n >= 7 && m < 7) {
        Hs4[1 * 49 + m * 7 + (n - 7)] = val;
        Hs4[2 * 49 + (n - 7) * 7 + m] = val;
      } else {
        Hs4[3 * 49 + (n - 7) * 7 + (m - 7)] = val;
        Hs4[3 * 49 + (m - 7) * 7 + (n - 7)] = val;
      }
      l++;
    }
  }
}

/* Per-edge blocks in the reference layout Hs [4,E,7,7], gs [2,E,7] (fp64).
 * ii_edge / jj_edge are ROWS into Twc/Xs/Cs (already searchsorted). */
ORACLE_API void oracle_edge_blocks(int mode, const float* Twc, const float* Xs, const float* Cs,
                                   const float* Kmat /*[3,3] or NULL*/, const int64_t* ii_edge,
                                   const int64_t* jj_edge, const int64_t* idx /*[E,N]*/,
                                   const uint8_t* valid /*[E,N]*/, const float* Q /*[E,N]*/,
                                   int K, int N, int E, int height, int width, int pixel_border,
                                   float z_eps, float sigma_a, float sigma_b, float C_thresh,
                                   float Q_thresh, double* Hs, double* gs) {
  (void)K;
  gn_params_t P;
  memset(&P, 0, sizeof P);
  P.mode = mode;
  P.sigma_a = sigma_a;
  P.sigma_b = sigma_b;
  P.C_thresh = C_thresh;
  P.Q_thresh = Q_thresh;
  if (Kmat) {
    P.fx = Kmat[0];
    P.fy = Kmat[4];
    P.cx = Kmat[2];
    P.cy = Kmat[5];
  }
  P.height = height;
  P.width = width;
  P.pixel_border = pixel_border;
  P.z_eps = z_eps;
#pragma omp parallel for schedule(dynamic, 1)
  for (int e = 0; e < E; e++) {
    double Hs4[4 * 49], gs2[14];
    edge_blocks(&P, Twc, Xs, Cs, (int)ii_edge[e], (int)jj_edge[e], idx + (size_t)e * N,
                valid + (size_t)e * N, Q + (size_t)e * N, N, Hs4, gs2);
    for (int b = 0; b < 4; b++)
      memcpy(Hs + ((size_t)b * E + e) * 49, Hs4 + b * 49, 49 * sizeof(double));
    for (int b = 0; b < 2; b++) memcpy(gs + ((size_t)b * E + e) * 7, gs2 + b * 7, 7 * sizeof(double));
  }
}

/* ------------------------------------------------------------------------ */
/* index helpers (gn_kernels.cu:203-212): unique (sorted) + searchsorted     */
/* ------------------------------------------------------------------------ */

static int cmp_i64(const void* a, const void* b) {
  int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
  return (x > y) - (x < y);
}

/* returns number of unique ids; ii_edge/jj_edge = rank of ii/jj among them */
ORACLE_API int oracle_edge_index(const int64_t* ii, const int64_t* jj, int E, int64_t* ii_edge,
                                 int64_t* jj_edge) {
  int64_t* all = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)(E > 0 ? E : 1));
  for (int e = 0; e < E; e++) {
    all[e] = ii[e];
    all[E + e] = jj[e];
  }
  qsort(all, 2 * (size_t)E, sizeof(int64_t), cmp_i64);
  int nu = 0;
  for (int k = 0; k < 2 * E; k++)
    if (k == 0 || all[k] != all[k - 1]) all[nu++] = all[k];
  for (int e = 0; e < E; e++) {
    int64_t key = ii[e];
    int lo = 0, hi = nu;
    while (lo < hi) { int mid = (lo + hi) / 2; if (all[mid] < key) lo = mid + 1; else hi = mid; }
    ii_edge[e] = lo;
    key = jj[e];
    lo = 0; hi = nu;
    while (lo < hi) { int mid = (lo + hi) / 2; if (all[mid] < key) lo = mid + 1; else hi = mid; }
    jj_edge[e] = lo;
  }
  free(all);
  return nu;
}

/* ------------------------------------------------------------------------ */
/* assembly (SparseBlock::update_lhs / update_rhs, gn_kernels.cu:72-122)     */
/* and solve (SparseBlock::solve :163-199 : LLT, failure -> zeros)           */
/* ------------------------------------------------------------------------ */

/* Hs fp64 [4,E,7,7], gs fp64 [2,E,7] -> dense H [n,n] row-major, g [n], n = 7*(K-num_fix).
 * The reference rounds the per-edge blocks to fp32 before assembly; pass
 * round_f32 = 1 to mimic (Hs produced by its fp32 kernel). */
ORACLE_API void oracle_assemble(const double* Hs, const double* gs, const int64_t* ii_edge,
                                const int64_t* jj_edge, int E, int K, int num_fix, int round_f32,
                                double* H, double* g) {
  const int n = 7 * (K - num_fix);
  memset(H, 0, sizeof(double) * (size_t)n * n);
  memset(g, 0, sizeof(double) * (size_t)n);
  for (int blk = 0; blk < 4; blk++) {
    for (int e = 0; e < E; e++) {
      const int64_t io = ii_edge[e] - num_fix, jo = jj_edge[e] - num_fix;
      const int64_t r = (blk < 2) ? io : jo;
      const int64_t c = (blk % 2 == 0) ? io : jo;
      if (r < 0 || c < 0) continue;
      const double* src = Hs + ((size_t)blk * E + e) * 49;
      for (int k = 0; k < 7; k++)
        for (int l = 0; l < 7; l++) {
          double val = round_f32 ? (double)(float)src[k * 7 + l] : src[k * 7 + l];
          H[(size_t)(7 * r + k) * n + (7 * c + l)] += val;
        }
    }
  }
  for (int blk = 0; blk < 2; blk++) {
    for (int e = 0; e < E; e++) {
      const int64_t r = ((blk == 0) ? ii_edge[e] : jj_edge[e]) - num_fix;
      if (r < 0) continue;
      const double* src = gs + ((size_t)blk * E + e) * 7;
      for (int k = 0; k < 7; k++) g[7 * r + k] += round_f32 ? (double)(float)src[k] : src[k];
    }
  }
}

/* Dense fp64 Cholesky solve H x = g (lower triangle of H is read, as Eigen's
 * SimplicialLLT does).  Returns 0 on success, 1 if H is not positive definite
 * (then x is all zeros, gn_kernels.cu:193-197). */
ORACLE_API int oracle_chol_solve(const double* H, const double* g, int n, double* x) {
  double* L = (double*)malloc(sizeof(double) * (size_t)n * n);
  memcpy(L, H, sizeof(double) * (size_t)n * n);
  int fail = 0;
  for (int j = 0; j < n && !fail; j++) {
    double d = L[(size_t)j * n + j];
    for (int k = 0; k < j; k++) d -= L[(size_t)j * n + k] * L[(size_t)j * n + k];
    if (!(d > 0.0)) {
      fail = 1;
      break;
    }
    d = sqrt(d);
    L[(size_t)j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double s = L[(size_t)i * n + j];
      for (int k = 0; k < j; k++) s -= L[(size_t)i * n + k] * L[(size_t)j * n + k];
      L[(size_t)i * n + j] = s / d;
    }
  }
  if (fail) {
    memset(x, 0, sizeof(double) * (size_t)n);
    free(L);
    return 1;
  }
  for (int i = 0; i < n; i++) { /* L y = g */
    double s = g[i];
    for (int k = 0; k < i; k++) s -= L[(size_t)i * n + k] * x[k];
    x[i] = s / L[(size_t)i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) { /* L^T x = y */
    double s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[(size_t)k * n + i] * x[k];
    x[i] = s / L[(size_t)i * n + i];
  }
  free(L);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* full drivers: gauss_newton_{rays,points,calib}_cuda                       */
/*   gn_kernels.cu:909-995, :2206-2293, :2610-2702                           */
/* ------------------------------------------------------------------------ */

/* mode 0 rays / 1 points / 2 calib.  Twc is updated in place.  dx_out
 * [K-num_fix,7] receives the last iteration's update.  H_first/g_first (may be
 * NULL) receive the first iteration's assembled system (test aid).
 * Returns the number of iterations executed. */
/* ------------------------------------------------------------------------------------------
 * gauss_newton_rays_odom extras (gn_kernels.cu:1792-1934): odometry constraints
 * (odom_constraint_kernel_left_perturb_log, :1106-1286, with its helpers quat_to_rot :562-574,
 * vec_to_skew / R_to_vec / Jl_inv / logSE3 :1000-1086) restated in fp32, one odometry edge
 * at a time, and the scale prior (apply_scale_prior_cpu, :1716-1790) in fp64.
 * ------------------------------------------------------------------------------------------ */
/* fp32 expressions below carry the FMA contractions nvcc 12.9 -O3 applies to the reference
 * source (read from the SASS): a*b + c*d -> fma(a, b, round(c*d)); x +/- a*b -> fma; loops
 * `acc += a*b` -> fma chains.  They matter: the rotation residual goes through
 * acosf(0.5 (trace - 1)), which quantises angles below ~5e-4 rad, so one ulp in R_err can move
 * the odometry residual by 3e-4 rad at an information of 1e6. */
static void quat_to_rot_o(const float* q, float* R) { /* :562-574 */
  const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
  const float x2 = qx + qx, y2 = qy + qy, z2 = qz + qz;
  const float wz = qw * z2, wy = qw * y2, wx = qw * x2;
  const float ax = fmaf(-qx, x2, 1.0f), ay = fmaf(-qy, y2, 1.0f);
  R[0] = fmaf(-qz, z2, ay);
  R[1] = fmaf(qy, x2, -wz);
  R[2] = fmaf(qz, x2, wy);
  R[3] = fmaf(qy, x2, wz);
  R[4] = fmaf(-qz, z2, ax);
  R[5] = fmaf(qz, y2, -wx);
  R[6] = fmaf(qz, x2, -wy);
  R[7] = fmaf(qz, y2, wx);
  R[8] = fmaf(-qy, y2, ax);
}

static void skew_o(const float* v, float* M) { /* :1000-1004 */
  M[0] = 0;     M[1] = -v[2]; M[2] = v[1];
  M[3] = v[2];  M[4] = 0;     M[5] = -v[0];
  M[6] = -v[1]; M[7] = v[0];  M[8] = 0;
}

static void mat33_vec3(const float* A, const float* x, float* y) { /* :1088-1092 */
  y[0] = fmaf(A[2], x[2], fmaf(A[0], x[0], A[1] * x[1]));
  y[1] = fmaf(A[5], x[2], fmaf(A[3], x[0], A[4] * x[1]));
  y[2] = fmaf(A[8], x[2], fmaf(A[6], x[0], A[7] * x[1]));
}

static void mat33_mat33(const float* A, const float* B, float* C) { /* :1095-1104 */
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      float acc = 0;
      for (int k = 0; k < 3; ++k) acc = fmaf(A[r * 3 + k], B[k * 3 + c], acc);
      C[r * 3 + c] = acc;
    }
}

static void transpose33(const float* A, float* T) {
  T[0] = A[0]; T[1] = A[3]; T[2] = A[6];
  T[3] = A[1]; T[4] = A[4]; T[5] = A[7];
  T[6] = A[2]; T[7] = A[5]; T[8] = A[8];
}

static void R_to_vec_o(const float* R, float* omega) { /* :1007-1041 */
  const float trace = R[0] + R[4] + R[8];
  float cos_angle = 0.5f * (trace - 1.0f);
  cos_angle = fmaxf(-1.0f, fminf(1.0f, cos_angle));
  const float angle = acosf(cos_angle);
  if (fabsf(angle) < 1e-6f) {
    omega[0] = omega[1] = omega[2] = 0.0f;
  } else if (fabsf((float)(angle - M_PI)) < 1e-6f) { /* M_PI is a double: fp64 difference */
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

static void Jl_inv_o(const float* omega, float* J_inv) { /* :1044-1074 */
  const float angle =
      sqrtf(fmaf(omega[2], omega[2], fmaf(omega[0], omega[0], omega[1] * omega[1])));
  float sk[9];
  skew_o(omega, sk);
  for (int i = 0; i < 9; ++i) J_inv[i] = (i % 4 == 0) ? 1.0f : 0.0f;
  if (fabsf(angle) < 1e-6f) {
    for (int i = 0; i < 9; ++i) J_inv[i] = fmaf(-0.5f, sk[i], J_inv[i]);
  } else {
    const float angle_sq = angle * angle;
    const float sin_angle = sinf(angle), cos_angle = cosf(angle);
    const float factor1 = -0.5f;
    const float factor2 = (1.0f / angle_sq) - ((1.0f + cos_angle) / (2.0f * angle * sin_angle));
    float sk2[9];
    mat33_mat33(sk, sk, sk2);
    for (int i = 0; i < 9; ++i) J_inv[i] += fmaf(factor1, sk[i], factor2 * sk2[i]);
  }
}

/* Hs_odom [4,E,7,7], gs_odom [2,E,7] (fp32, reference layout).  odom_ii / odom_jj are used
 * RAW as rows of Twc (:1121-1122); Tij rows hold [t(3), q_xyzw(4), ...] with `stride` floats. */
ORACLE_API void oracle_odom_blocks(const float* Twc, const float* Tij, int stride,
                                   const int64_t* odom_ii, const int64_t* odom_jj, int Eo,
                                   float sigma_odom_t, float sigma_odom_r, float* Hs, float* gs) {
  for (int idx = 0; idx < Eo; ++idx) {
    const float* Ti = Twc + 8 * odom_ii[idx];
    const float* Tj = Twc + 8 * odom_jj[idx];
    const float *ti = Ti, *qi = Ti + 3, *tj = Tj, *qj = Tj + 3;
    const float sj = Tj[7];
    const float* tm = Tij + (size_t)stride * idx;
    float R_m[9], R_i[9], R_j[9], R_i_T[9], R_j_T[9], R_m_T[9];
    quat_to_rot_o(tm + 3, R_m);
    quat_to_rot_o(qi, R_i);
    quat_to_rot_o(qj, R_j);
    transpose33(R_i, R_i_T);
    transpose33(R_j, R_j_T);
    transpose33(R_m, R_m_T);
    float R_ij[9];
    mat33_mat33(R_i_T, R_j, R_ij);
    const float t_diff[3] = {tj[0] - ti[0], tj[1] - ti[1], tj[2] - ti[2]};
    float R_iT_tdiff[3];
    mat33_vec3(R_i_T, t_diff, R_iT_tdiff);
    const float sj_inv = (sj == 0.0f) ? 1e6f : 1.0f / sj;
    const float t_pred[3] = {sj_inv * R_iT_tdiff[0], sj_inv * R_iT_tdiff[1], sj_inv * R_iT_tdiff[2]};
    float t_meas_inv[3];
    mat33_vec3(R_m_T, tm, t_meas_inv);
    t_meas_inv[0] *= -1.0f; t_meas_inv[1] *= -1.0f; t_meas_inv[2] *= -1.0f;
    float R_err[9], Rm_T_tpred[3], t_err[3];
    mat33_mat33(R_m_T, R_ij, R_err);
    mat33_vec3(R_m_T, t_pred, Rm_T_tpred);
    for (int k = 0; k < 3; ++k) t_err[k] = Rm_T_tpred[k] + t_meas_inv[k];
    float r[6], J_inv[9];
    R_to_vec_o(R_err, r + 3);
    Jl_inv_o(r + 3, J_inv);
    mat33_vec3(J_inv, t_err, r);
    float Wd[6];
    const float info_t = 1.0f / (sigma_odom_t * sigma_odom_t);
    const float info_r = 1.0f / (sigma_odom_r * sigma_odom_r);
    for (int k = 0; k < 3; ++k) Wd[k] = info_t * huber(r[k] / sigma_odom_t);
    for (int k = 3; k < 6; ++k) Wd[k] = info_r * huber(r[k] / sigma_odom_r);
    float J_i[42] = {0}, J_j[42] = {0};
    float tj_inv[3], tj_sk[9], drho[9], drho_ds[3];
    mat33_vec3(R_j_T, tj, tj_inv);
    tj_inv[0] *= -1.0; tj_inv[1] *= -1.0; tj_inv[2] *= -1.0; /* fp64 literal, exact */
    skew_o(tj_inv, tj_sk);
    mat33_mat33(tj_sk, R_j_T, drho);
    mat33_vec3(R_m_T, t_pred, drho_ds);
    for (int rr = 0; rr < 3; ++rr) {
      for (int c = 0; c < 3; ++c) {
        J_j[rr * 7 + c] = R_j_T[rr * 3 + c];            /* d rho / d t_j */
        J_j[rr * 7 + 3 + c] = drho[rr * 3 + c];         /* d rho / d theta_j */
        J_j[(rr + 3) * 7 + 3 + c] = R_j_T[rr * 3 + c];  /* d phi / d theta_j */
      }
      J_j[rr * 7 + 6] = -drho_ds[rr];                   /* d rho / d s_j */
    }
    for (int k = 0; k < 42; ++k) J_i[k] = -J_j[k];
    J_i[6] = J_i[13] = J_i[20] = 0.0f;
    J_i[27] = J_i[34] = J_i[41] = 0.0f;
    float J[84], JTW[84];
    for (int rr = 0; rr < 6; ++rr)
      for (int c = 0; c < 7; ++c) {
        J[rr * 14 + c] = J_i[rr * 7 + c];
        J[rr * 14 + 7 + c] = J_j[rr * 7 + c];
      }
    for (int a2 = 0; a2 < 14; ++a2)
      for (int k = 0; k < 6; ++k) JTW[a2 * 6 + k] = J[k * 14 + a2] * Wd[k];
    float Hf[196], gf[14];
    for (int a2 = 0; a2 < 14; ++a2) {
      for (int b2 = 0; b2 < 14; ++b2) {
        float val = 0;
        for (int k = 0; k < 6; ++k) val = fmaf(JTW[a2 * 6 + k], J[k * 14 + b2], val);
        Hf[a2 * 14 + b2] = val;
      }
      float val = 0;
      for (int k = 0; k < 6; ++k) val = fmaf(JTW[a2 * 6 + k], r[k], val);
      gf[a2] = val;
    }
    for (int n2 = 0; n2 < 7; ++n2) {
      gs[(0 * (size_t)Eo + idx) * 7 + n2] = gf[n2];
      gs[(1 * (size_t)Eo + idx) * 7 + n2] = gf[n2 + 7];
    }
    for (int rr = 0; rr < 7; ++rr)
      for (int c = 0; c < 7; ++c) {
        Hs[((0 * (size_t)Eo + idx) * 7 + rr) * 7 + c] = Hf[rr * 14 + c];
        Hs[((1 * (size_t)Eo + idx) * 7 + rr) * 7 + c] = Hf[rr * 14 + c + 7];
        Hs[((2 * (size_t)Eo + idx) * 7 + rr) * 7 + c] = Hf[(rr + 7) * 14 + c];
        Hs[((3 * (size_t)Eo + idx) * 7 + rr) * 7 + c] = Hf[(rr + 7) * 14 + c + 7];
      }
  }
}

ORACLE_API int oracle_gauss_newton(int mode, float* Twc, const float* Xs, const float* Cs,
                                   const float* Kmat, const int64_t* ii, const int64_t* jj,
                                   const int64_t* idx, const uint8_t* valid, const float* Q, int K,
                                   int N, int E, int height, int width, int pixel_border,
                                   float z_eps, float sigma_a, float sigma_b, float C_thresh,
                                   float Q_thresh, int num_fix, int max_iter, float delta_thresh,
                                   float* dx_out, double* H_first, double* g_first,
                                   int* solve_fail_count) {
  const int n = 7 * (K - num_fix);
  int64_t* ii_edge = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E > 0 ? E : 1));
  int64_t* jj_edge = (int64_t*)malloc(sizeof(int64_t) * (size_t)(E > 0 ? E : 1));
  oracle_edge_index(ii, jj, E, ii_edge, jj_edge);
  double* Hs = (double*)malloc(sizeof(double) * 4 * 49 * (size_t)(E > 0 ? E : 1));
  double* gs = (double*)malloc(sizeof(double) * 2 * 7 * (size_t)(E > 0 ? E : 1));
  double* H = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1) * (n > 0 ? n : 1));
  double* g = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  double* x = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  int iters = 0, fails = 0;
  for (int itr = 0; itr < max_iter; itr++) {
    oracle_edge_blocks(mode, Twc, Xs, Cs, Kmat, ii_edge, jj_edge, idx, valid, Q, K, N, E, height,
                       width, pixel_border, z_eps, sigma_a, sigma_b, C_thresh, Q_thresh, Hs, gs);
    oracle_assemble(Hs, gs, ii_edge, jj_edge, E, K, num_fix, 1, H, g);
    if (itr == 0) {
      if (H_first) memcpy(H_first, H, sizeof(double) * (size_t)n * n);
      if (g_first) memcpy(g_first, g, sizeof(double) * (size_t)n);
    }
    fails += oracle_chol_solve(H, g, n, x);
    double nrm = 0.0;
    for (int k = 0; k < n; k++) {
      float d = (float)(-x[k]); /* dx = -A.solve(), fp64 -> fp32 (:977, :191-192) */
      dx_out[k] = d;
      nrm += (double)d * (double)d;
    }
    oracle_pose_retr(Twc, dx_out, K, num_fix);
    iters++;
    if ((float)sqrt(nrm) < delta_thresh) break; /* :986-989 */
  }
  if (solve_fail_count) *solve_fail_count = fails;
  free(ii_edge);
  free(jj_edge);
  free(Hs);
  free(gs);
  free(H);
  free(g);
  free(x);
  return iters;
}
