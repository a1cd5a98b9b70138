// C ABI of libfrodo_b200.so (see include/frodo_b200.h).
#include <stdarg.h>
#include <string.h>

#include "gn_internal.cuh"

namespace frodo {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return FRODO_ERR_CUDA;
}

int num_sms() {
  static int cache[kMaxDevices] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return kNumSMs;
  if (cache[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kNumSMs;
    cache[dev] = n;
  }
  return cache[dev];
}

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

size_t gn_carve_workspace(void* base, int K, int E, int N, GnWorkspace* ws) {
  const size_t Ez = (size_t)(E > 0 ? E : 1);
  const size_t n = (size_t)7 * (size_t)(K > 0 ? K : 1);
  const size_t max_chunks = ((size_t)(N > 0 ? N : 1) + 511) / 512;  // >= tiles per edge
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes, 256);
    return o;
  };
  // K-dependent regions first: frodo_gn_solve_retract locates them without knowing E, N
  const size_t o_ch = take(gn_chol_scratch_bytes((int)n));
  const size_t o_hg = take((n + 1) * n * sizeof(double));
  const size_t o_ii = take(Ez * sizeof(int32_t));
  const size_t o_jj = take(Ez * sizeof(int32_t));
  const size_t o_sc = take((8 + (size_t)(K > 0 ? K : 1)) * sizeof(int32_t));
  const size_t o_tk = take(Ez * sizeof(int32_t));
  const size_t o_pt = take(Ez * max_chunks * kGnBlock * sizeof(float));
  const size_t o_eb = take(Ez * kGnBlock * sizeof(double));
  // last: the pointmaps repacked in two layouts (frodo_gn_pack_points), ~32 bytes per point
  const size_t o_pk = take(gn_packed_bytes(K, N));
  if (base != nullptr && ws != nullptr) {
    char* b = static_cast<char*>(base);
    ws->chol = reinterpret_cast<double*>(b + o_ch);
    ws->Hg = reinterpret_cast<double*>(b + o_hg);
    ws->ii_row = reinterpret_cast<int32_t*>(b + o_ii);
    ws->jj_row = reinterpret_cast<int32_t*>(b + o_jj);
    ws->scalars = reinterpret_cast<int32_t*>(b + o_sc);
    ws->kf_need = ws->scalars + 8;
    ws->tickets = reinterpret_cast<int32_t*>(b + o_tk);
    ws->partials = reinterpret_cast<float*>(b + o_pt);
    ws->edge_blocks = reinterpret_cast<double*>(b + o_eb);
    ws->packed = reinterpret_cast<float*>(b + o_pk);
    ws->bytes = off;
  }
  if (ws != nullptr) ws->packed_off = o_pk;
  return off;
}

}  // namespace frodo

using namespace frodo;

extern "C" {

int frodo_version(void) { return 100; }

const char* frodo_last_error(void) { return g_err; }

int frodo_iter_proj(const float* rays_img, const float* pts_3d_norm, const float* p_init,
                    float* p_new, uint8_t* converged, int B, int H, int W, int N, int max_iter,
                    float lambda_init, float cost_thresh, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && N >= 0 && H >= 3 && W >= 3);
  FRODO_ARG_CHECK(B == 0 || N == 0 ||
                  (rays_img && pts_3d_norm && p_init && p_new && converged));
  return launch_iter_proj(rays_img, pts_3d_norm, p_init, p_new, converged, B, H, W, N, max_iter,
                          lambda_init, cost_thresh, static_cast<cudaStream_t>(stream));
}

int frodo_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1, int64_t* p1_new,
                             void* max_scores, int B, int H, int W, int N, int F, int radius,
                             int dilation_max, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && N >= 0 && H >= 1 && W >= 1 && F >= 1 && radius >= 0);
  FRODO_ARG_CHECK(B == 0 || N == 0 || (D11 && D21 && p1 && p1_new && max_scores));
  return launch_refine_f16(static_cast<const __half*>(D11), static_cast<const __half*>(D21), p1,
                           p1_new, static_cast<__half*>(max_scores), B, H, W, N, F, radius,
                           dilation_max, static_cast<cudaStream_t>(stream));
}

int frodo_refine_matches_f64(const double* D11, const double* D21, const int64_t* p1,
                             int64_t* p1_new, double* max_scores, int B, int H, int W, int N,
                             int F, int radius, int dilation_max, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && N >= 0 && H >= 1 && W >= 1 && F >= 1 && radius >= 0);
  FRODO_ARG_CHECK(B == 0 || N == 0 || (D11 && D21 && p1 && p1_new && max_scores));
  return launch_refine_f64(D11, D21, p1, p1_new, max_scores, B, H, W, N, F, radius, dilation_max,
                           static_cast<cudaStream_t>(stream));
}

int frodo_set_refine_patch(int patch) { return refine_forced_patch(patch < 0 ? 0 : patch); }

int frodo_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                             int64_t* p1_new, float* max_scores, int B, int H, int W, int N, int F,
                             int radius, int dilation_max, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && N >= 0 && H >= 1 && W >= 1 && F >= 1 && radius >= 0);
  FRODO_ARG_CHECK(B == 0 || N == 0 || (D11 && D21 && p1 && p1_new && max_scores));
  return launch_refine_f32(D11, D21, p1, p1_new, max_scores, B, H, W, N, F, radius, dilation_max,
                           static_cast<cudaStream_t>(stream));
}

// ---- the matching pipeline around the kernels (matching.py:25-110) ----------------------
int frodo_match_prep(const float* X11, const float* X21, const int64_t* idx_1_to_2_init,
                     float* rays_with_grad, float* pts3d_norm, float* p_init, int B, int H, int W,
                     frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && H >= 2 && W >= 2);
  FRODO_ARG_CHECK(B == 0 || (X11 && X21 && rays_with_grad && pts3d_norm && p_init));
  return launch_match_prep(X11, X21, idx_1_to_2_init, rays_with_grad, pts3d_norm, p_init, B, H, W,
                           static_cast<cudaStream_t>(stream));
}

int frodo_match_occlusion(const float* X11, const float* X21, const float* p_iter,
                          const uint8_t* converged, float dist_thresh, int64_t* p1, uint8_t* valid,
                          int B, int H, int W, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && H >= 1 && W >= 1);
  FRODO_ARG_CHECK(B == 0 || (X11 && X21 && p_iter && converged && p1 && valid));
  return launch_match_occlusion(X11, X21, p_iter, converged, dist_thresh, p1, valid, B, H, W,
                                static_cast<cudaStream_t>(stream));
}

int frodo_convert_f32_f16(const float* src, void* dst, size_t n, frodo_stream_t stream) {
  FRODO_ARG_CHECK(n == 0 || (src && dst));
  return launch_convert_f32_f16(src, static_cast<__half*>(dst), n,
                                static_cast<cudaStream_t>(stream));
}

int frodo_constrain_points_to_ray(const float* Xs, const float* Kmat, int num_frames, int H, int W,
                                  float* Xs_out, frodo_stream_t stream) {
  FRODO_ARG_CHECK(num_frames >= 0 && H >= 1 && W >= 1);
  FRODO_ARG_CHECK(num_frames == 0 || (Xs && Kmat && Xs_out));
  return launch_constrain_points(Xs, Kmat, num_frames, H, W, Xs_out,
                                 static_cast<cudaStream_t>(stream));
}

int frodo_match_finish(const int64_t* p1, const void* scores, int scores_f16,
                       const uint8_t* valid_in, float threshold, int64_t* idx_1_to_2,
                       uint8_t* valid, int B, int N, int W, frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && N >= 0 && W >= 1);
  FRODO_ARG_CHECK(B == 0 || N == 0 || (p1 && scores && valid_in && idx_1_to_2 && valid));
  return launch_match_finish(p1, scores, scores_f16 != 0, valid_in, threshold, idx_1_to_2, valid,
                             B, N, W, static_cast<cudaStream_t>(stream));
}

namespace {
// Workspace of frodo_match, carved in this order (each piece 256-byte aligned).
struct MatchCarve {
  size_t rays, pts, p_init, p_iter, conv, p1, valid, d11h, d21h, p1_new, scores, total;
};
MatchCarve match_carve(int B, int H, int W, int F) {
  const size_t px = (size_t)(B > 0 ? B : 0) * (size_t)H * (size_t)W;
  MatchCarve c;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t at = off;
    off += (bytes + 255) & ~(size_t)255;
    return at;
  };
  c.rays = take(px * 36);
  c.pts = take(px * 12);
  c.p_init = take(px * 8);
  c.p_iter = take(px * 8);
  c.conv = take(px);
  c.p1 = take(px * 16);
  c.valid = take(px);
  c.d11h = take(px * (size_t)F * 2);
  c.d21h = take(px * (size_t)F * 2);
  c.p1_new = take(px * 16);
  c.scores = take(px * 2);
  c.total = off;
  return c;
}
}  // namespace

size_t frodo_match_workspace_bytes(int B, int H, int W, int F) {
  if (B < 0 || H < 1 || W < 1 || F < 1) return 0;
  return match_carve(B, H, W, F).total;
}

int frodo_match(const frodo_match_params* prm, const float* X11, const float* X21,
                const float* D11, const float* D21, const int64_t* idx_1_to_2_init,
                int64_t* idx_1_to_2, uint8_t* valid, int B, int H, int W, int F, void* workspace,
                size_t workspace_bytes, frodo_stream_t stream) {
  FRODO_ARG_CHECK(prm != nullptr);
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && H >= 3 && W >= 3 && F >= 1);
  FRODO_ARG_CHECK(prm->radius >= 1 && prm->dilation_max >= 1 && prm->max_iter >= 0);
  if (B == 0) return FRODO_OK;
  FRODO_ARG_CHECK(X11 && X21 && D11 && D21 && idx_1_to_2 && valid && workspace);
  const MatchCarve c = match_carve(B, H, W, F);
  FRODO_ARG_CHECK(workspace_bytes >= c.total);
  FRODO_ARG_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255) == 0);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  float* rays = reinterpret_cast<float*>(ws + c.rays);
  float* pts = reinterpret_cast<float*>(ws + c.pts);
  float* p_init = reinterpret_cast<float*>(ws + c.p_init);
  float* p_iter = reinterpret_cast<float*>(ws + c.p_iter);
  uint8_t* conv = ws + c.conv;
  int64_t* p1 = reinterpret_cast<int64_t*>(ws + c.p1);
  uint8_t* v0 = ws + c.valid;
  __half* d11h = reinterpret_cast<__half*>(ws + c.d11h);
  __half* d21h = reinterpret_cast<__half*>(ws + c.d21h);
  int64_t* p1_new = reinterpret_cast<int64_t*>(ws + c.p1_new);
  __half* scores = reinterpret_cast<__half*>(ws + c.scores);
  const int N = H * W;
  const size_t nd = (size_t)B * N * F;
  int rc;
  if ((rc = launch_match_prep(X11, X21, idx_1_to_2_init, rays, pts, p_init, B, H, W, s))) return rc;
  // iter_proj with the occlusion test (matching.py:66-75) in its epilogue: p_iter / converged are
  // never written, p1 and the first validity mask come straight from the registers
  const IterProjTail itail = {X11, X21, prm->dist_thresh, p1, v0};
  if ((rc = launch_iter_proj(rays, pts, p_init, p_iter, conv, B, H, W, N, prm->max_iter,
                             prm->lambda_init, prm->convergence_thresh, s, &itail)))
    return rc;
  if ((rc = launch_convert_f32_f16(D11, d11h, nd, s))) return rc;
  // the tile-staged refine kernel converts the query descriptors while loading them
  // (matching.py:80 fused): the cast pass over D21 and its fp16 copy are only needed otherwise
  const bool fuse_d21 = refine_f16_takes_tile_kernel(d11h, D21, p1, p1_new, B, H, W, N, F);
  if (!fuse_d21 && (rc = launch_convert_f32_f16(D21, d21h, nd, s))) return rc;
  if (fuse_d21) {
    // ... and the threshold + linearisation (matching.py:87-92) in refine_matches' epilogue:
    // four kernels per call (prep, iter_proj, cast of D11, refine_matches)
    const RefineTail rtail = {v0, prm->ft_matching_threshold, idx_1_to_2, valid};
    return launch_refine_f16(d11h, d21h, p1, p1_new, scores, B, H, W, N, F, prm->radius,
                             prm->dilation_max, s, D21, &rtail);
  }
  if ((rc = launch_refine_f16(d11h, d21h, p1, p1_new, scores, B, H, W, N, F, prm->radius,
                              prm->dilation_max, s, nullptr)))
    return rc;
  return launch_match_finish(p1_new, scores, true, v0, prm->ft_matching_threshold, idx_1_to_2,
                             valid, B, N, W, s);
}

size_t frodo_gn_workspace_bytes(int K, int E, int N) {
  return gn_carve_workspace(nullptr, K, E, N, nullptr);
}

int frodo_gn_edge_index(const int64_t* ii, const int64_t* jj, int E, int32_t* ii_row,
                        int32_t* jj_row, int32_t* n_unique, frodo_stream_t stream) {
  FRODO_ARG_CHECK(E >= 0);
  FRODO_ARG_CHECK(E == 0 || (ii && jj && ii_row && jj_row));
  return launch_gn_edge_index(ii, jj, E, ii_row, jj_row, n_unique,
                              static_cast<cudaStream_t>(stream));
}

size_t frodo_gn_workspace_packed_offset(int K, int E, int N) {
  GnWorkspace ws;
  gn_carve_workspace(nullptr, K, E, N, &ws);
  return ws.packed_off;
}

size_t frodo_gn_packed_bytes(int K, int N) { return gn_packed_bytes(K, N); }

int frodo_gn_pack_points(const float* Xs, const float* Cs, int K, int N, float* XC,
                         frodo_stream_t stream) {
  FRODO_ARG_CHECK(K >= 0 && N >= 0);
  FRODO_ARG_CHECK(K == 0 || N == 0 || (Xs && Cs && XC));
  FRODO_ARG_CHECK((reinterpret_cast<uintptr_t>(XC) & 15) == 0);
  return launch_gn_pack_points(Xs, Cs, K, N, XC, static_cast<cudaStream_t>(stream));
}

int frodo_gn_accumulate(const frodo_gn_params* params, const float* Twc, const float* Xs,
                        const float* Cs, const float* Kmat, const float* XC,
                        const int32_t* ii_row,
                        const int32_t* jj_row, const int64_t* idx_ii2jj,
                        const uint8_t* valid_match, const float* Q, int K, int N, int E,
                        int num_fix, double* Hg, double* edge_blocks, const int32_t* done,
                        void* workspace, size_t workspace_bytes, frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr && K >= 1 && N >= 0 && E >= 0 && num_fix >= 0);
  FRODO_ARG_CHECK(params->mode != FRODO_GN_CALIB || Kmat != nullptr);
  FRODO_ARG_CHECK(E <= 65535);
  FRODO_ARG_CHECK(workspace != nullptr);
  FRODO_ARG_CHECK(workspace_bytes >= gn_carve_workspace(nullptr, K, E, N, nullptr));
  FRODO_ARG_CHECK(E == 0 || (Twc && Xs && Cs && ii_row && jj_row && idx_ii2jj && valid_match && Q));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GnWorkspace ws;
  gn_carve_workspace(workspace, K, E, N, &ws);
  double* eb = edge_blocks ? edge_blocks : ws.edge_blocks;
  int rc = launch_gn_accumulate(*params, Twc, Xs, Cs, Kmat, XC, ii_row, jj_row, idx_ii2jj,
                                valid_match, Q, K, N, E, ws.partials, ws.tickets, eb, done, s);
  if (rc) return rc;
  if (Hg == nullptr) return FRODO_OK;  // per-edge blocks only
  return launch_gn_assemble(eb, ii_row, jj_row, E, K, num_fix, Hg, done, s);
}

int frodo_gn_solve_retract(double* Hg, float* Twc, float* dx, int K, int num_fix,
                           float delta_thresh, int32_t* done, int32_t* info, void* workspace,
                           size_t workspace_bytes, frodo_stream_t stream) {
  FRODO_ARG_CHECK(K >= 1 && num_fix >= 0 && num_fix <= K);
  FRODO_ARG_CHECK(K == num_fix || (Hg && Twc && dx));
  // the Cholesky tile scratch is the first region of the workspace (depends on K only)
  double* chol = nullptr;
  if (workspace != nullptr && workspace_bytes >= gn_chol_scratch_bytes(7 * K))
    chol = static_cast<double*>(workspace);
  const GnSystem sys = {Hg, true, nullptr, nullptr, nullptr, 0};
  return launch_gn_solve_retract(sys, Twc, dx, K, num_fix, delta_thresh, done, info, chol, 7 * K,
                                 static_cast<cudaStream_t>(stream));
}

size_t frodo_gn_chol_stamps_offset(int K) { return gn_chol_stamps_offset(7 * (K > 0 ? K : 1)); }

int frodo_gn_workspace_init(void* workspace, size_t workspace_bytes, int K, int E, int N,
                            frodo_stream_t stream) {
  FRODO_ARG_CHECK(workspace != nullptr);
  FRODO_ARG_CHECK(workspace_bytes >= gn_carve_workspace(nullptr, K, E, N, nullptr));
  GnWorkspace ws;
  gn_carve_workspace(workspace, K, E, N, &ws);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  FRODO_CUDA_TRY(cudaMemsetAsync(ws.scalars, 0, 8 * sizeof(int32_t), s));
  FRODO_CUDA_TRY(cudaMemsetAsync(ws.tickets, 0, (size_t)(E > 0 ? E : 1) * sizeof(int32_t), s));
  return gn_chol_scratch_init(ws.chol, 7 * (K > 0 ? K : 1), s);
}

// Shared driver of the single-call Gauss-Newton entry points.  `od` (may be NULL) adds the
// odometry / scale-prior terms of gauss_newton_rays_odom; its scratch follows the base workspace.
static int gn_run(const frodo_gn_params* params, float* Twc, const float* Xs, const float* Cs,
                  const float* Kmat, const int64_t* ii, const int64_t* jj,
                  const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q, int K,
                  int N, int E, int num_fix, int max_iter, float delta_thresh, float* dx,
                  int32_t* info, void* workspace, size_t workspace_bytes, const GnOdom* od,
                  frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr && K >= 1 && N >= 0 && E >= 0 && num_fix >= 0 && num_fix <= K);
  FRODO_ARG_CHECK(workspace != nullptr);
  const size_t base_bytes = gn_carve_workspace(nullptr, K, E, N, nullptr);
  FRODO_ARG_CHECK(workspace_bytes >= base_bytes + (od ? gn_odom_scratch_bytes(od->Eo) : 0));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GnWorkspace ws;
  gn_carve_workspace(workspace, K, E, N, &ws);
  void* odom_scratch = static_cast<char*>(workspace) + base_bytes;  // base_bytes is 256-aligned
  int rc = frodo_gn_workspace_init(workspace, workspace_bytes, K, E, N, stream);
  if (rc) return rc;
  int32_t* done = ws.scalars + 1;
  int32_t* winfo = ws.scalars + 2;
  if (od != nullptr && od->Eo > 0) {
    // ranks of the odometry endpoints against the visual keyframe list: ii ranks, then jj ranks
    int32_t* rows = static_cast<int32_t*>(odom_scratch);
    rc = launch_gn_edge_index(ii, jj, E, ws.ii_row, ws.jj_row, ws.scalars, s, od->odom_ii, od->Eo,
                              rows);
    if (rc) return rc;
    rc = launch_gn_edge_index(ii, jj, E, ws.ii_row, ws.jj_row, ws.scalars, s, od->odom_jj, od->Eo,
                              rows + od->Eo);
  } else {
    rc = launch_gn_edge_index(ii, jj, E, ws.ii_row, ws.jj_row, ws.scalars, s);
  }
  if (rc) return rc;
  // packed fast path: repack the pointmaps once per solve (they do not change across iterations)
  const float* XC = nullptr;
  if (E > 0 && N > 0 && (N % 16) == 0 && max_iter > 0) {
    rc = launch_gn_pack_points(Xs, Cs, K, N, ws.packed, s);
    if (rc) return rc;
    XC = ws.packed;
  }
  for (int it = 0; it < max_iter; ++it) {
    // the visual terms stay per-edge blocks (the solver builds its tiles from them); only the
    // odometry terms need the dense system
    rc = frodo_gn_accumulate(params, Twc, Xs, Cs, Kmat, XC, ws.ii_row, ws.jj_row, idx_ii2jj,
                             valid_match, Q, K, N, E, num_fix, od != nullptr ? ws.Hg : nullptr,
                             ws.edge_blocks, done, workspace, workspace_bytes, stream);
    if (rc) return rc;
    if (od != nullptr) {
      rc = launch_gn_add_odom(*od, Twc, K, num_fix, ws.Hg, done, odom_scratch, s);
      if (rc) return rc;
    }
    const GnSystem sys = {ws.Hg, od != nullptr, ws.edge_blocks, ws.ii_row, ws.jj_row, E};
    rc = launch_gn_solve_retract(sys, Twc, dx, K, num_fix, delta_thresh, done, winfo, ws.chol,
                                 7 * K, s);
    if (rc) return rc;
  }
  if (info != nullptr)
    FRODO_CUDA_TRY(cudaMemcpyAsync(info, winfo, 2 * sizeof(int32_t), cudaMemcpyDeviceToDevice, s));
  return FRODO_OK;
}

int frodo_gauss_newton(const frodo_gn_params* params, float* Twc, const float* Xs, const float* Cs,
                       const float* Kmat, const int64_t* ii, const int64_t* jj,
                       const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q, int K,
                       int N, int E, int num_fix, int max_iter, float delta_thresh, float* dx,
                       int32_t* info, void* workspace, size_t workspace_bytes,
                       frodo_stream_t stream) {
  return gn_run(params, Twc, Xs, Cs, Kmat, ii, jj, idx_ii2jj, valid_match, Q, K, N, E, num_fix,
                max_iter, delta_thresh, dx, info, workspace, workspace_bytes, nullptr, stream);
}

size_t frodo_gn_odom_workspace_bytes(int K, int E, int N, int E_odom) {
  return gn_carve_workspace(nullptr, K, E, N, nullptr) + gn_odom_scratch_bytes(E_odom);
}

int frodo_gauss_newton_rays_odom(const frodo_gn_params* params, float* Twc, const float* Xs,
                                 const float* Cs, const int64_t* ii, const int64_t* jj,
                                 const int64_t* idx_ii2jj, const uint8_t* valid_match,
                                 const float* Q, const int64_t* odom_ii, const int64_t* odom_jj,
                                 const float* Tij, int tij_stride, int E_odom, const double* s_bar,
                                 float sigma_odom_t, float sigma_odom_r, float sigma_scale_prior,
                                 int K, int N, int E, int num_fix, int max_iter,
                                 float delta_thresh, float* dx, int32_t* info, void* workspace,
                                 size_t workspace_bytes, frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr && params->mode == FRODO_GN_RAYS);
  FRODO_ARG_CHECK(E_odom >= 0 && (E_odom == 0 || (odom_ii && odom_jj && Tij && tij_stride >= 7)));
  GnOdom od;
  od.odom_ii = odom_ii;
  od.odom_jj = odom_jj;
  od.Tij = Tij;
  od.stride = tij_stride;
  od.Eo = E_odom;
  od.s_bar = s_bar;
  od.sigma_t = sigma_odom_t;
  od.sigma_r = sigma_odom_r;
  od.sigma_scale = sigma_scale_prior;
  return gn_run(params, Twc, Xs, Cs, nullptr, ii, jj, idx_ii2jj, valid_match, Q, K, N, E, num_fix,
                max_iter, delta_thresh, dx, info, workspace, workspace_bytes, &od, stream);
}

size_t frodo_gn_odom_blocks_bytes(int E_odom) { return gn_odom_blocks_bytes(E_odom); }

static GnOdom odom_from_terms(const frodo_gn_odom_terms* t) {
  GnOdom od;
  od.odom_ii = t->odom_ii_row;
  od.odom_jj = t->odom_jj_row;
  od.Tij = t->Tij;
  od.stride = t->tij_stride;
  od.Eo = t->E_odom;
  od.s_bar = t->s_bar;
  od.sigma_t = t->sigma_odom_t;
  od.sigma_r = t->sigma_odom_r;
  od.sigma_scale = t->sigma_scale_prior;
  return od;
}

static int check_odom_terms(const frodo_gn_odom_terms* t) {
  FRODO_ARG_CHECK(t != nullptr && t->E_odom >= 0);
  FRODO_ARG_CHECK(t->E_odom == 0 || (t->odom_ii_row && t->odom_jj_row && t->odom_rank && t->Tij &&
                                     t->tij_stride >= 7));
  FRODO_ARG_CHECK(t->blocks != nullptr && t->blocks_bytes >= gn_odom_blocks_bytes(t->E_odom));
  return FRODO_OK;
}

int frodo_gn_add_odom(const frodo_gn_odom_terms* terms, const float* Twc,
                      const int32_t* pose_rows, int K, int num_fix, double* Hg,
                      const int32_t* done, frodo_stream_t stream) {
  FRODO_ARG_CHECK(K >= 1 && num_fix >= 0 && num_fix <= K);
  FRODO_ARG_CHECK(K == num_fix || (Twc && Hg));
  int rc = check_odom_terms(terms);
  if (rc) return rc;
  const GnOdom od = odom_from_terms(terms);
  return launch_gn_add_odom_rows(od, Twc, K, num_fix, Hg, done, terms->odom_rank,
                                 static_cast<double*>(terms->blocks), pose_rows,
                                 static_cast<cudaStream_t>(stream));
}

/* ---- device-resident keyframe / edge arena -------------------------------- */
int frodo_gn_pack_slot(const float* X, const float* C, int K_arena, int slot, int N, float* XC,
                       frodo_stream_t stream) {
  FRODO_ARG_CHECK(K_arena >= 1 && slot >= 0 && slot < K_arena && N >= 0);
  FRODO_ARG_CHECK(N == 0 || (X && C && XC));
  FRODO_ARG_CHECK((reinterpret_cast<uintptr_t>(XC) & 15) == 0);
  return launch_gn_pack_slot(X, C, K_arena, slot, N, XC, static_cast<cudaStream_t>(stream));
}

size_t frodo_gn_arena_workspace_bytes(int K_arena, int E, int N) {
  GnWorkspace ws;
  gn_carve_workspace(nullptr, K_arena, E, N, &ws);
  return ws.packed_off;  // everything but the packed pointmaps, which the arena owns
}

int frodo_gauss_newton_arena(const frodo_gn_params* params, float* Twc, const float* Xs,
                             const float* Cs, const float* Kmat, const float* XC_packed,
                             const int32_t* ii_slot, const int32_t* jj_slot,
                             const int32_t* ii_rank, const int32_t* jj_rank,
                             const int32_t* pose_slot, const int64_t* idx_ii2jj,
                             const uint8_t* valid_match, const float* Q, int K_arena, int K_solve,
                             int N, int E, int num_fix, int max_iter, float delta_thresh,
                             const frodo_gn_odom_terms* odom, float* dx, int32_t* info,
                             void* workspace, size_t workspace_bytes, frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr && K_arena >= 1 && K_solve >= 1 && K_solve <= K_arena);
  if (odom != nullptr) {
    FRODO_ARG_CHECK(params->mode == FRODO_GN_RAYS);
    const int orc = check_odom_terms(odom);
    if (orc) return orc;
  }
  FRODO_ARG_CHECK(N >= 0 && E >= 0 && E <= 65535 && num_fix >= 0 && num_fix <= K_solve);
  FRODO_ARG_CHECK(params->mode != FRODO_GN_CALIB || Kmat != nullptr);
  FRODO_ARG_CHECK(workspace != nullptr &&
                  workspace_bytes >= frodo_gn_arena_workspace_bytes(K_arena, E, N));
  FRODO_ARG_CHECK(Twc && pose_slot && (K_solve == num_fix || dx));
  FRODO_ARG_CHECK(E == 0 || (Xs && Cs && ii_slot && jj_slot && ii_rank && jj_rank && idx_ii2jj &&
                             valid_match && Q));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GnWorkspace ws;
  gn_carve_workspace(workspace, K_arena, E, N, &ws);
  FRODO_CUDA_TRY(cudaMemsetAsync(ws.scalars, 0, 8 * sizeof(int32_t), s));
  FRODO_CUDA_TRY(cudaMemsetAsync(ws.tickets, 0, (size_t)(E > 0 ? E : 1) * sizeof(int32_t), s));
  {
    const int irc = gn_chol_scratch_init(ws.chol, 7 * K_arena, s);
    if (irc) return irc;
  }
  int32_t* done = ws.scalars + 1;
  int32_t* winfo = ws.scalars + 2;
  const float* XC = (N % 16) == 0 ? XC_packed : nullptr;
  for (int it = 0; it < max_iter; ++it) {
    // data rows = arena slots; system rows = ranks among the participating keyframes
    int rc = launch_gn_accumulate(*params, Twc, Xs, Cs, Kmat, XC, ii_slot, jj_slot, idx_ii2jj,
                                  valid_match, Q, K_arena, N, E, ws.partials, ws.tickets,
                                  ws.edge_blocks, done, s);
    if (rc) return rc;
    if (odom != nullptr) {
      rc = launch_gn_assemble(ws.edge_blocks, ii_rank, jj_rank, E, K_solve, num_fix, ws.Hg, done,
                              s);
      if (rc) return rc;
      const GnOdom od = odom_from_terms(odom);
      rc = launch_gn_add_odom_rows(od, Twc, K_solve, num_fix, ws.Hg, done, odom->odom_rank,
                                   static_cast<double*>(odom->blocks), pose_slot, s);
      if (rc) return rc;
    }
    const GnSystem sys = {ws.Hg, odom != nullptr, ws.edge_blocks, ii_rank, jj_rank, E};
    rc = launch_gn_solve_retract(sys, Twc, dx, K_solve, num_fix, delta_thresh, done, winfo,
                                 ws.chol, 7 * K_arena, s, pose_slot);
    if (rc) return rc;
  }
  if (info != nullptr)
    FRODO_CUDA_TRY(cudaMemcpyAsync(info, winfo, 2 * sizeof(int32_t), cudaMemcpyDeviceToDevice, s));
  return FRODO_OK;
}

/* ---- edge-sharded multi-GPU driver: the whole solve in one call -------------- */
size_t frodo_gn_peer_buffer_bytes(int max_edges) {
  return (size_t)2 * (size_t)(max_edges > 0 ? max_edges : 1) * kGnBlock * sizeof(double) + 256;
}

int frodo_gauss_newton_sharded(const frodo_gn_params* params, float* Twc, const float* Xs,
                               const float* Cs, const float* Kmat, const int32_t* ii_row,
                               const int32_t* jj_row, int E, int edge_lo, int edge_hi,
                               const int64_t* idx_local, const uint8_t* valid_local,
                               const float* Q_local, int K, int N, int num_fix, int max_iter,
                               float delta_thresh, const frodo_gn_peer* peer,
                               const frodo_gn_odom_terms* odom, float* dx, int32_t* done,
                               int32_t* info, void* workspace, size_t workspace_bytes,
                               frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr && K >= 1 && N >= 0 && E >= 0 && num_fix >= 0 && num_fix <= K);
  FRODO_ARG_CHECK(edge_lo >= 0 && edge_lo <= edge_hi && edge_hi <= E && E <= 65535);
  FRODO_ARG_CHECK(params->mode != FRODO_GN_CALIB || Kmat != nullptr);
  const int El = edge_hi - edge_lo;
  FRODO_ARG_CHECK(workspace != nullptr && done != nullptr);
  FRODO_ARG_CHECK(workspace_bytes >= gn_carve_workspace(nullptr, K, E > 0 ? E : 1, N, nullptr));
  FRODO_ARG_CHECK(Twc && (K == num_fix || dx));
  FRODO_ARG_CHECK(E == 0 || (Xs && Cs && ii_row && jj_row));
  FRODO_ARG_CHECK(El == 0 || (idx_local && valid_local && Q_local));
  const bool multi = peer != nullptr && peer->world > 1;
  if (multi) {
    FRODO_ARG_CHECK(peer->world <= 16 && peer->rank >= 0 && peer->rank < peer->world);
    FRODO_ARG_CHECK(peer->peer_buffers != nullptr && peer->local_buffer != nullptr);
    FRODO_ARG_CHECK(peer->max_edges >= E);
  }
  if (odom != nullptr) {
    FRODO_ARG_CHECK(params->mode == FRODO_GN_RAYS);
    const int orc = check_odom_terms(odom);
    if (orc) return orc;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  GnWorkspace ws;
  gn_carve_workspace(workspace, K, E > 0 ? E : 1, N, &ws);  // sized for all E edges
  int rc = frodo_gn_workspace_init(workspace, workspace_bytes, K, E > 0 ? E : 1, N, stream);
  if (rc) return rc;
  const float* XC = nullptr;
  if (El > 0 && N > 0 && (N % 16) == 0 && max_iter > 0) {
    const int32_t* need = nullptr;
    if (El < E) {  // a slice of the edges reads only some of the pointmaps
      rc = launch_gn_mark_keyframes(ii_row + edge_lo, jj_row + edge_lo, El, K, ws.kf_need, s);
      if (rc) return rc;
      need = ws.kf_need;
    }
    rc = launch_gn_pack_points(Xs, Cs, K, N, ws.packed, s, need);
    if (rc) return rc;
    XC = ws.packed;
  }
  // layout of a rank's exchange buffer: [blocks of even epochs | blocks of odd epochs | counters]
  const size_t eb_bytes = multi ? (size_t)peer->max_edges * kGnBlock * sizeof(double) : 0;
  uint32_t arrivals[2] = {multi ? peer->arrivals0[0] : 0u, multi ? peer->arrivals0[1] : 0u};
  for (int it = 0; it < max_iter; ++it) {
    const double* eb_all = ws.edge_blocks;
    const uint32_t* wait_counter = nullptr;
    uint32_t wait_target = 0;
    GnPeerOut po = {};
    if (multi) {
      const uint32_t epoch = peer->epoch0 + (uint32_t)it + 1u;
      const int par = (int)(epoch & 1u);
      po.bufs = static_cast<void* const*>(peer->peer_buffers);
      po.eb_offset = (size_t)par * eb_bytes;
      po.cnt_offset = 2 * eb_bytes + (size_t)par * 128;
      po.world = peer->world;
      po.edge_base = edge_lo;
      char* local = static_cast<char*>(peer->local_buffer);
      eb_all = reinterpret_cast<const double*>(local + po.eb_offset);
      wait_counter = reinterpret_cast<const uint32_t*>(local + po.cnt_offset);
      arrivals[par] += (uint32_t)E;  // every epoch brings E arrivals, on every rank
      wait_target = arrivals[par];
    }
    rc = launch_gn_accumulate(*params, Twc, Xs, Cs, Kmat, XC, ii_row + edge_lo, jj_row + edge_lo,
                              idx_local, valid_local, Q_local, K, N, El, ws.partials, ws.tickets,
                              ws.edge_blocks + (size_t)edge_lo * kGnBlock, done, s,
                              multi ? &po : nullptr);
    if (rc) return rc;
    GnSystem sys = {ws.Hg, false, eb_all, ii_row, jj_row, E};
    sys.wait_counter = wait_counter;
    sys.wait_target = wait_target;
    if (odom != nullptr) {  // the odometry terms need the dense system (replicated, O(K) work)
      if (wait_counter != nullptr) {
        rc = launch_gn_peer_wait(wait_counter, wait_target, done, info, s);
        if (rc) return rc;
      }
      rc = launch_gn_assemble(eb_all, ii_row, jj_row, E, K, num_fix, ws.Hg, done, s);
      if (rc) return rc;
      const GnOdom od = odom_from_terms(odom);
      rc = launch_gn_add_odom_rows(od, Twc, K, num_fix, ws.Hg, done, odom->odom_rank,
                                   static_cast<double*>(odom->blocks), nullptr, s);
      if (rc) return rc;
      sys.dense = true;
      sys.wait_counter = nullptr;
    }
    rc = launch_gn_solve_retract(sys, Twc, dx, K, num_fix, delta_thresh, done, info, ws.chol,
                                 7 * K, s);
    if (rc) return rc;
  }
  return FRODO_OK;
}

int frodo_gn_peer_allreduce(const void* peer_buffers, const void* peer_signal_pads,
                            size_t buffer_offset_bytes, int signal_slot0, int rank, int world,
                            uint32_t epoch, double* out, size_t n, const int32_t* done,
                            frodo_stream_t stream) {
  FRODO_ARG_CHECK(peer_buffers != nullptr && peer_signal_pads != nullptr && out != nullptr);
  FRODO_ARG_CHECK(world >= 1 && world <= 16 && rank >= 0 && rank < world && signal_slot0 >= 0);
  FRODO_ARG_CHECK((buffer_offset_bytes & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0);
  FRODO_ARG_CHECK(epoch != 0);
  return launch_peer_allreduce(peer_buffers, peer_signal_pads, buffer_offset_bytes, signal_slot0,
                               rank, world, epoch, out, n, done, static_cast<cudaStream_t>(stream));
}

int frodo_match_quality(const float* Q_src, const float* Q_dst, const float* C_src,
                        const float* C_dst, const int64_t* idx, const uint8_t* valid_match,
                        float Q_conf, float C_conf, float* Q_out, uint8_t* valid_opt,
                        uint8_t* valid_q, int32_t* counts, int B, int N, int N_src,
                        frodo_stream_t stream) {
  FRODO_ARG_CHECK(B >= 0 && B <= 65535 && N >= 0 && N_src >= 0);
  FRODO_ARG_CHECK(B == 0 || counts != nullptr);
  FRODO_ARG_CHECK(B == 0 || N == 0 ||
                  (Q_src && Q_dst && idx && valid_match && Q_out && valid_opt && valid_q &&
                   N_src > 0));
  return launch_match_quality(Q_src, Q_dst, C_src, C_dst, idx, valid_match, Q_conf, C_conf, Q_out,
                              valid_opt, valid_q, counts, B, N, N_src,
                              static_cast<cudaStream_t>(stream));
}

size_t frodo_track_workspace_bytes(int N) { return track_workspace_bytes(N); }

int frodo_track_pose(const frodo_track_params* params, const float* Xf, const int64_t* idx_f2k,
                     const float* Xk, const float* Qk, const uint8_t* valid, const float* meas_k,
                     const uint8_t* valid_meas_k, const float* Kmat, const float* T_WCf,
                     const float* T_WCk, int N, float* T_WCf_new, float* T_CkCf, int32_t* info,
                     double* costs, void* workspace, size_t workspace_bytes,
                     frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr);
  FRODO_ARG_CHECK(params->mode == FRODO_GN_RAYS || params->mode == FRODO_GN_CALIB);
  FRODO_ARG_CHECK(params->sigma_a > 0.0f && params->sigma_b > 0.0f && params->huber > 0.0f);
  FRODO_ARG_CHECK(params->max_iters >= 0 && N >= 0);
  FRODO_ARG_CHECK(T_WCf && T_WCk && T_WCf_new && T_CkCf);
  FRODO_ARG_CHECK(N == 0 || (Xf && Qk && valid));
  if (params->mode == FRODO_GN_RAYS) {
    FRODO_ARG_CHECK(N == 0 || Xk != nullptr);
  } else {
    FRODO_ARG_CHECK(Kmat != nullptr);
    FRODO_ARG_CHECK(N == 0 || (params->raw_calib ? Xk != nullptr : (meas_k && valid_meas_k)));
    FRODO_ARG_CHECK(params->height > 0 && params->width > 0);
    FRODO_ARG_CHECK(!params->raw_calib || (long long)params->height * params->width == N);
  }
  FRODO_ARG_CHECK(workspace != nullptr && workspace_bytes >= track_workspace_bytes(N));
  FRODO_ARG_CHECK((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0);
  return launch_track_pose(*params, Xf, idx_f2k, Xk, Qk, valid, meas_k, valid_meas_k, Kmat, T_WCf,
                           T_WCk, N, T_WCf_new, T_CkCf, info, costs, workspace,
                           static_cast<cudaStream_t>(stream));
}

int frodo_track_pose_static_max_points(void) { return track_static_max_points(); }

int frodo_track_pose_static(const frodo_track_params* params, const float* Xf,
                            const int64_t* idx_f2k, const float* Xk, const float* Qk,
                            const uint8_t* valid, const float* meas_k,
                            const uint8_t* valid_meas_k, const float* Kmat, const float* T_WCf,
                            const float* T_WCk, int N, float* T_WCf_new, float* T_CkCf,
                            int32_t* info, double* costs, frodo_stream_t stream) {
  FRODO_ARG_CHECK(params != nullptr);
  FRODO_ARG_CHECK(params->mode == FRODO_GN_RAYS || params->mode == FRODO_GN_CALIB);
  FRODO_ARG_CHECK(params->sigma_a > 0.0f && params->sigma_b > 0.0f && params->huber > 0.0f);
  FRODO_ARG_CHECK(params->max_iters >= 0 && N >= 0);
  FRODO_ARG_CHECK(T_WCf && T_WCk && T_WCf_new && T_CkCf);
  FRODO_ARG_CHECK(N == 0 || (Xf && Qk && valid));
  if (params->mode == FRODO_GN_RAYS) {
    FRODO_ARG_CHECK(N == 0 || Xk != nullptr);
  } else {
    FRODO_ARG_CHECK(Kmat != nullptr);
    FRODO_ARG_CHECK(N == 0 || (params->raw_calib ? Xk != nullptr : (meas_k && valid_meas_k)));
    FRODO_ARG_CHECK(params->height > 0 && params->width > 0);
    FRODO_ARG_CHECK(!params->raw_calib || (long long)params->height * params->width == N);
  }
  return launch_track_pose_static(*params, Xf, idx_f2k, Xk, Qk, valid, meas_k, valid_meas_k, Kmat,
                                  T_WCf, T_WCk, N, T_WCf_new, T_CkCf, info, costs,
                                  static_cast<cudaStream_t>(stream));
}

}  // extern "C"
