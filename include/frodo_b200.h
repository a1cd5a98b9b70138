/*
 * frodo_b200.h -- C ABI of libfrodo_b200.so, the B200 (sm_100a) implementation
 * of the reference's `mast3r_slam_backends` operator surface.
 *
 * Every entry point takes raw DEVICE pointers, plain sizes and scalars, and a
 * CUDA stream handle (cudaStream_t passed as void*; NULL = legacy default
 * stream).  All calls are asynchronous with respect to the host unless noted.
 * Return value: 0 = ok, FRODO_ERR_ARG = bad argument, FRODO_ERR_CUDA = a CUDA
 * runtime call or launch failed (text via frodo_last_error()).  A failed
 * Cholesky (system not positive definite) is NOT an error: dx is zero and the
 * poses are left unchanged, as in the reference (gn_kernels.cu:193-197).
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   mast3r_slam/backend/src/gn.cpp:120-135   iter_proj
 *   mast3r_slam/backend/src/gn.cpp:137-150   refine_matches
 *   mast3r_slam/backend/src/gn.cpp:63-88     gauss_newton_rays
 *   mast3r_slam/backend/src/gn.cpp:3-26      gauss_newton_points
 *   mast3r_slam/backend/src/gn.cpp:90-118    gauss_newton_calib
 *   mast3r_slam/backend/include/gn.h         (*_cuda launchers)
 *   mast3r_slam/backend/src/gn.cpp:29-62     gauss_newton_rays_odom
 *   mast3r_slam/matching.py:8-110            match() and its helpers (Python in the reference)
 *   mast3r_slam/tracker.py:397-517           the tracker's pose solve (Python in the reference)
 *   mast3r_slam/global_opt.py:186-224        per-solve graph packing (Python in the reference)
 */
#ifndef FRODO_B200_H
#define FRODO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FRODO_OK 0
#define FRODO_ERR_ARG 1
#define FRODO_ERR_CUDA 2

#define FRODO_GN_RAYS 0
#define FRODO_GN_POINTS 1
#define FRODO_GN_CALIB 2

typedef void* frodo_stream_t; /* cudaStream_t */

/* Library version (major*10000 + minor*100 + patch). */
int frodo_version(void);
/* Text of the last error on the calling thread ("" if none). */
const char* frodo_last_error(void);

/* ---- iter_proj: matching_kernels.cu:128-325 -------------------------------
 * rays_img [B,H,W,9] f32 (ray, d ray/dx, d ray/dy), pts_3d_norm [B,N,3] f32,
 * p_init [B,N,2] f32 (u,v).  Outputs p_new [B,N,2] f32, converged [B,N] u8
 * (0/1, layout-compatible with torch.bool).  Bit-exact with the reference
 * kernel as compiled by nvcc 12.9 -O3 (same FMA contractions). */
int frodo_iter_proj(const float* rays_img, const float* pts_3d_norm, const float* p_init,
                    float* p_new, uint8_t* converged, int B, int H, int W, int N, int max_iter,
                    float lambda_init, float cost_thresh, frodo_stream_t stream);

/* ---- refine_matches: matching_kernels.cu:25-125 ---------------------------
 * D11 [B,H,W,F], D21 [B,N,F] (IEEE fp16 for _f16, fp32 for _f32, fp64 for _f64: the three
 * dispatch cases of the reference, matching_kernels.cu:112), p1 [B,N,2]
 * i64 (u,v).  Outputs p1_new [B,N,2] i64, max_scores [B,N] (same dtype as D).
 * fp16 path rounds to fp16 after every multiply and every add (c10::Half). */
int frodo_refine_matches_f16(const void* D11, const void* D21, const int64_t* p1, int64_t* p1_new,
                             void* max_scores, int B, int H, int W, int N, int F, int radius,
                             int dilation_max, frodo_stream_t stream);
int frodo_refine_matches_f32(const float* D11, const float* D21, const int64_t* p1,
                             int64_t* p1_new, float* max_scores, int B, int H, int W, int N, int F,
                             int radius, int dilation_max, frodo_stream_t stream);
int frodo_refine_matches_f64(const double* D11, const double* D21, const int64_t* p1,
                             int64_t* p1_new, double* max_scores, int B, int H, int W, int N,
                             int F, int radius, int dilation_max, frodo_stream_t stream);
/* Tuning / test switch of the tile-staged fp16 kernel: query-patch edge 16 or 32, 0 = automatic
 * (16 for F <= 24).  Both variants are bit-identical; returns the value now in force.  The
 * environment variable FRODO_REFINE_PATCH seeds it once per process. */
int frodo_set_refine_patch(int patch);

/* ---- the matching pipeline around the two kernels ----------
 * mast3r_slam/matching.py:25-110 + mast3r_slam/image.py:5-38, which the reference runs as
 * ~20 PyTorch kernels per call.  The arithmetic order is fixed, every product and sum
 * rounded separately (no FMA): |x|^2 = (x0 x0 + x1 x1) + x2 x2, IEEE sqrt and division,
 * x / max(|x|, 1e-12); Scharr taps (3/32, 10/32) accumulated in row-major order, zero taps
 * skipped.  torch orders the same sums differently, so against the reference the float
 * outputs agree to 1-2 ulp and the final indices / masks are identical on the golden
 * vectors (tests/golden/golden_match.npz); against the CPU restatement under oracle/ (same
 * order) every output is bit-identical.
 *
 * frodo_match_prep: prep_for_iter_proj (matching.py:25-49).  X11, X21 [B,H,W,3] f32;
 *   idx_1_to_2_init [B,H*W] i64 or NULL (identity).  Outputs rays_with_grad [B,H,W,9]
 *   (normalised ray, Scharr d/dx, d/dy with reflect padding), pts3d_norm [B,H*W,3],
 *   p_init [B,H*W,2] f32 (x, y).  H, W >= 2.
 * frodo_match_occlusion: matching.py:66-75.  p1 [B,H*W,2] i64 = trunc(p_iter);
 *   valid = converged & (||X11[p1] - X21|| < dist_thresh).
 * frodo_convert_f32_f16: tensor.half() (matching.py:79-80), round to nearest even.
 * frodo_match_finish: matching.py:87-92.  valid = valid_in & (score > threshold, compared
 *   in the scores' dtype), idx_1_to_2 = x + W * y.  scores_f16 != 0: scores are fp16.
 * frodo_match: match() (matching.py:8-10) = all of the above + iter_proj + refine_matches
 *   (fp16 descriptors) in one call; D11 [B,H,W,F], D21 [B,H,W,F] f32; outputs idx_1_to_2
 *   [B,H*W] i64 and valid [B,H*W] u8; workspace of frodo_match_workspace_bytes(B,H,W,F). */
typedef struct frodo_match_params {
  int max_iter;
  float lambda_init;
  float convergence_thresh;
  float dist_thresh;
  int radius;       /* 0: no refinement is not supported by match() in the reference either */
  int dilation_max;
  float ft_matching_threshold;
} frodo_match_params;

int frodo_match_prep(const float* X11, const float* X21, const int64_t* idx_1_to_2_init,
                     float* rays_with_grad, float* pts3d_norm, float* p_init, int B, int H, int W,
                     frodo_stream_t stream);
int frodo_match_occlusion(const float* X11, const float* X21, const float* p_iter,
                          const uint8_t* converged, float dist_thresh, int64_t* p1, uint8_t* valid,
                          int B, int H, int W, frodo_stream_t stream);
int frodo_convert_f32_f16(const float* src, void* dst, size_t n, frodo_stream_t stream);
/* constrain_points_to_ray (mast3r_slam/geometry.py:37-42, :107-115), the re-projection
 * solve_GN_calib applies to the pointmaps before gauss_newton_calib (global_opt.py:296-300):
 * Xs [num_frames, H*W, 3] f32 -> Xs_out (may alias Xs) = z * ((u - cx)/fx, (v - cy)/fy, 1) with
 * (u, v) the pixel of the point and Kmat [3,3] f32 on the device.  IEEE operations in the
 * reference's order: bit-identical to the torch expression. */
int frodo_constrain_points_to_ray(const float* Xs, const float* Kmat, int num_frames, int H, int W,
                                  float* Xs_out, frodo_stream_t stream);
int frodo_match_finish(const int64_t* p1, const void* scores, int scores_f16,
                       const uint8_t* valid_in, float threshold, int64_t* idx_1_to_2,
                       uint8_t* valid, int B, int N, int W, frodo_stream_t stream);
size_t frodo_match_workspace_bytes(int B, int H, int W, int F);
int frodo_match(const frodo_match_params* params, const float* X11, const float* X21,
                const float* D11, const float* D21, const int64_t* idx_1_to_2_init,
                int64_t* idx_1_to_2, uint8_t* valid, int B, int H, int W, int F, void* workspace,
                size_t workspace_bytes, frodo_stream_t stream);

/* frodo_match_quality: what the callers compute from a match before using it -- the tracker
 *   (mast3r_slam/tracker.py:194-206) and FactorGraph.add_factors (global_opt.py:98-112):
 *     Q[b,n]         = sqrt(Q_src[b, idx[b,n]] * Q_dst[b,n])
 *     valid_q[b,n]   = valid_match[b,n] & (Q[b,n] > Q_conf)                 (valid_kf / valid_j)
 *     valid_opt[b,n] = valid_q & (C_src[b, idx[b,n]] > C_conf) & (C_dst[b,n] > C_conf)
 *     counts[b]      = {sum valid_opt, sum valid_q}  (match_frac = counts / N), int32 [B,2]
 *   Q_src, C_src [B,N_src] f32 (C_src / C_dst may be NULL: no confidence test), Q_dst, C_dst
 *   [B,N] f32, idx [B,N] i64 in [0,N_src), valid_match [B,N] u8.  Bit-identical to torch. */
int frodo_match_quality(const float* Q_src, const float* Q_dst, const float* C_src,
                        const float* C_dst, const int64_t* idx, const uint8_t* valid_match,
                        float Q_conf, float C_conf, float* Q_out, uint8_t* valid_opt,
                        uint8_t* valid_q, int32_t* counts, int B, int N, int N_src,
                        frodo_stream_t stream);

/* ---- Gauss-Newton over the keyframe graph ---------------------------------
 * Shared tensor meanings (gn_kernels.cu:909-995):
 *   Twc [K,8] f32 = [t(3), q_xyzw(4), s], updated IN PLACE;  Xs [K,N,3] f32;
 *   Cs [K,N] f32;  ii,jj [E] i64 keyframe ids;  idx_ii2jj [E,N] i64;
 *   valid_match [E,N] u8;  Q [E,N] f32;  Kmat [3,3] f32 (calib only).
 * Parameters of the three residual models travel in frodo_gn_params. */
typedef struct frodo_gn_params {
  int mode;      /* FRODO_GN_RAYS / POINTS / CALIB                              */
  float sigma_a; /* rays: sigma_ray   points: sigma_point   calib: sigma_pixel  */
  float sigma_b; /* rays: sigma_dist  points: unused        calib: sigma_depth  */
  float C_thresh;
  float Q_thresh;
  int height; /* calib only (gn_kernels.cu:2307-2314) */
  int width;
  int pixel_border;
  float z_eps;
} frodo_gn_params;

/* Bytes of device workspace the GN entry points need for a graph of K
 * keyframes, E (local) directed edges and N points per keyframe. */
size_t frodo_gn_workspace_bytes(int K, int E, int N);

/* Zeroes the control words (tickets, convergence flag, counters, the ready flags of the
 * multi-CTA Cholesky) of a workspace; call once per solve before the first
 * frodo_gn_accumulate / frodo_gn_solve_retract on a fresh or reused buffer. */
int frodo_gn_workspace_init(void* workspace, size_t workspace_bytes, int K, int E, int N,
                            frodo_stream_t stream);
/* Profiling aid: byte offset (inside a workspace laid out for K keyframes) of the clock64
 * phase stamps the Cholesky's spine CTA writes when the environment has FRODO_CHOL_STAMPS=1
 * (profiles/chol_phases.py). */
size_t frodo_gn_chol_stamps_offset(int K);

/* Rays fast path: repacks the keyframe pointmaps Xs [K,N,3] and confidences Cs [K,N] into
 * XC, a 16-byte aligned region of frodo_gn_packed_bytes(K, N) bytes: first [K,N] float4
 * (x,y,z,c), so that the data-dependent gather of ray_align_kernel
 * (gn_kernels.cu:677-681,714) is one aligned 16-byte load per point, then the same data
 * interleaved per point pair (xA,xB,yA,yB,zA,zB,cA,cB; A, B = points p, p+256 of a 512-point
 * block, zero padded to whole blocks) for the streamed frame.  The pointmaps do not change
 * across Gauss-Newton iterations: pack once per solve and hand XC to frodo_gn_accumulate.
 * frodo_gn_workspace_packed_offset gives the byte offset of such a region inside a workspace
 * of frodo_gn_workspace_bytes. */
size_t frodo_gn_packed_bytes(int K, int N);
int frodo_gn_pack_points(const float* Xs, const float* Cs, int K, int N, float* XC,
                         frodo_stream_t stream);
size_t frodo_gn_workspace_packed_offset(int K, int E, int N);

/* get_unique_kf_idx + create_inds (gn_kernels.cu:203-212): writes, for every
 * edge, the row of its two keyframes among the sorted unique ids of cat(ii,jj)
 * (ii_row, jj_row: [E] i32) and the number of unique ids (n_unique: [1] i32). */
int frodo_gn_edge_index(const int64_t* ii, const int64_t* jj, int E, int32_t* ii_row,
                        int32_t* jj_row, int32_t* n_unique, frodo_stream_t stream);

/* One accumulation pass (ray_align / point_align / calib_proj kernels,
 * gn_kernels.cu:576-907, 1936-2204, 2295-2607) over the E edges given, followed
 * by assembly (SparseBlock::update_lhs/rhs, :72-122) into the dense fp64 system
 *   Hg = [ H (n x n, row-major) | g (n) ],  n = 7*(K - num_fix)
 * which is OVERWRITTEN (not accumulated into); Hg == NULL skips the assembly and
 * only produces edge_blocks.  Edge rows come from
 * frodo_gn_edge_index.  edge_blocks (may be NULL) receives per-edge [E,35] f64:
 * the 28 upper-triangular entries of H_jj (row-major) then g_j; by
 * construction H_ii = H_jj = -H_ij = -H_ji and g_i = -g_j.
 * XC_packed (may be NULL): output of frodo_gn_pack_points for these Xs, Cs; enables the
 * fast rays kernel (needs N % 16 == 0), otherwise Xs / Cs are read directly.
 * `done` (device int32, may be NULL): when non-zero at kernel time the pass is
 * skipped (convergence reached, see frodo_gn_solve_retract). */
int frodo_gn_accumulate(const frodo_gn_params* params, const float* Twc, const float* Xs,
                        const float* Cs, const float* Kmat, const float* XC_packed,
                        const int32_t* ii_row,
                        const int32_t* jj_row, const int64_t* idx_ii2jj,
                        const uint8_t* valid_match, const float* Q, int K, int N, int E,
                        int num_fix, double* Hg, double* edge_blocks, const int32_t* done,
                        void* workspace, size_t workspace_bytes, frodo_stream_t stream);

/* dx = -H^{-1} g by dense fp64 Cholesky (SparseBlock::solve, :163-199; zeros on
 * failure), dx -> f32 [K-num_fix,7]; Twc[k] <- Exp(dx[k-num_fix]) * Twc[k] for
 * k >= num_fix (pose_retr_kernel :522-560); if ||dx||_2 < delta_thresh sets
 * *done = 1 (:986-989).  Hg is destroyed.  Skipped when *done != 0 on entry.
 * info (device int32[2], may be NULL): [0] += 1 per executed iteration,
 * [1] += 1 per failed factorisation.
 * workspace: a buffer of frodo_gn_workspace_bytes(K, E, N) bytes prepared by
 * frodo_gn_workspace_init (the multi-CTA factorisation keeps its tile array and ready flags
 * at its start); NULL or too small selects a single-CTA factorisation. */
int frodo_gn_solve_retract(double* Hg, float* Twc, float* dx, int K, int num_fix,
                           float delta_thresh, int32_t* done, int32_t* info, void* workspace,
                           size_t workspace_bytes, frodo_stream_t stream);

/* Whole single-GPU driver = gauss_newton_{rays,points,calib}_cuda: edge index,
 * then max_iter x (accumulate, solve_retract) with a device-side convergence
 * flag; no host synchronisation.  dx [K-num_fix,7] f32 receives the last
 * executed iteration's update.  info as above (zeroed by this call). */
int frodo_gauss_newton(const frodo_gn_params* params, float* Twc, const float* Xs, const float* Cs,
                       const float* Kmat, const int64_t* ii, const int64_t* jj,
                       const int64_t* idx_ii2jj, const uint8_t* valid_match, const float* Q, int K,
                       int N, int E, int num_fix, int max_iter, float delta_thresh, float* dx,
                       int32_t* info, void* workspace, size_t workspace_bytes,
                       frodo_stream_t stream);

/* ---- gauss_newton_rays_odom: gn.cpp:29-62 / gn_kernels.cu:1792-1934 -----------------
 * gauss_newton_rays plus, every iteration, relative-pose (odometry) constraints
 * (odom_constraint_kernel_left_perturb_log, gn_kernels.cu:1106-1286) and a scale prior
 * (apply_scale_prior_cpu, :1716-1790), added to the normal equations before the solve.
 *   odom_ii, odom_jj [E_odom] int64: keyframe ids of edge i -> j.  As in the reference they
 *     index the rows of Twc directly and are ranked against the keyframes of the visual
 *     edges for the linear system.
 *   Tij [E_odom, tij_stride] f32: measured relative pose (t(3), q_xyzw(4), ...), tij_stride >= 7.
 *   s_bar [K] f64, device (the reference takes a host tensor; the Python operator uploads it):
 *     scale targets, entries <= 1e-6 are skipped; NULL disables the prior.
 * Workspace: frodo_gn_odom_workspace_bytes(K, E, N, E_odom) bytes. */
size_t frodo_gn_odom_workspace_bytes(int K, int E, int N, int E_odom);
int frodo_gauss_newton_rays_odom(const frodo_gn_params* params, float* Twc, const float* Xs,
                                 const float* Cs, const int64_t* ii, const int64_t* jj,
                                 const int64_t* idx_ii2jj, const uint8_t* valid_match,
                                 const float* Q, const int64_t* odom_ii, const int64_t* odom_jj,
                                 const float* Tij, int tij_stride, int E_odom, const double* s_bar,
                                 float sigma_odom_t, float sigma_odom_r, float sigma_scale_prior,
                                 int K, int N, int E, int num_fix, int max_iter,
                                 float delta_thresh, float* dx, int32_t* info, void* workspace,
                                 size_t workspace_bytes, frodo_stream_t stream);

/* The odometry / scale-prior terms of gauss_newton_rays_odom as a separate step, for drivers
 * that own the iteration loop (the edge-sharded multi-GPU driver adds them after its all-reduce;
 * the arena driver below takes them as an argument).  Adds the blocks of
 * odom_constraint_kernel_left_perturb_log and apply_scale_prior_cpu (gn_kernels.cu:1106-1286,
 * :1716-1790) to Hg [(n+1) x n], n = 7 (K - num_fix), in odometry-edge order. */
typedef struct frodo_gn_odom_terms {
  const int64_t* odom_ii_row; /* [E_odom] rows of Twc holding the two poses of each edge        */
  const int64_t* odom_jj_row;
  const int32_t* odom_rank;   /* [2 E_odom] rows of the linear system + num_fix (searchsorted of
                                 the ids in the visual keyframe list, :1829): ii ranks, jj ranks */
  const float* Tij;           /* [E_odom, tij_stride] measured relative poses                    */
  int tij_stride;
  int E_odom;
  const double* s_bar;        /* [K] device, by system rank; NULL disables the scale prior       */
  float sigma_odom_t, sigma_odom_r, sigma_scale_prior;
  void* blocks;               /* scratch, frodo_gn_odom_blocks_bytes(E_odom) bytes               */
  size_t blocks_bytes;
} frodo_gn_odom_terms;
size_t frodo_gn_odom_blocks_bytes(int E_odom);
/* pose_rows (may be NULL): system rank -> row of Twc (arena); NULL = identity. */
int frodo_gn_add_odom(const frodo_gn_odom_terms* terms, const float* Twc,
                      const int32_t* pose_rows, int K, int num_fix, double* Hg,
                      const int32_t* done, frodo_stream_t stream);

/* ---- the multi-GPU exchange step over NVLink peer memory ------------------------
 * Sum of the per-rank systems [H | g] (n doubles each) in ONE kernel: cross-rank barrier on
 * signal words + fixed-order (rank 0..world-1) sum with loads from the peers' memory, so every
 * rank gets bit-identical totals.  The reference has no multi-GPU path; the NCCL all-reduce this
 * replaces remains the fallback when no peer mapping is available.
 *   peer_buffers: DEVICE array of `world` pointers to the ranks' symmetric buffers (same layout on
 *     every rank; e.g. torch symmetric memory `buffer_ptrs_dev`); the partial system of this call
 *     starts buffer_offset_bytes into each of them.  Callers alternate between two offsets
 *     (double buffering by call parity): then this call's barrier is the only one needed.
 *   peer_signal_pads: DEVICE array of `world` pointers to uint32 signal words (zero-initialised
 *     once); words [signal_slot0, signal_slot0 + world) are used.
 *   epoch: strictly increasing (by one) per call, identical on all ranks, never 0.
 *   out [n] f64: local result.  `done` (may be NULL): when non-zero the call is skipped (it must
 *     hold the same value on all ranks: it does, the poses are bit-identical). */
/* The edge-sharded solve in ONE call (the reference has no multi-GPU path): every
 * rank passes the replicated poses / pointmaps, the keyframe ranks of ALL E edges (ii_row, jj_row
 * from frodo_gn_edge_index on the full edge list) and the data of its own contiguous slice
 * [edge_lo, edge_hi) (idx_local, valid_local, Q_local: [edge_hi - edge_lo, N]).  Per iteration:
 *   accumulate  this rank's edges; the last CTA of every edge stores the edge's 35 fp64 values
 *               into EVERY rank's copy of the global per-edge block array over NVLink peer
 *               memory and bumps that rank's arrival counter (release, system scope): the
 *               exchange is fused into the producer, 280 bytes per edge per rank;
 *   solve       the Cholesky's tile build waits until the local counter shows all E edges of
 *               this epoch (acquire), assembles in fixed edge order -- bit-identical systems on
 *               all ranks -- factors, retracts; every rank runs it redundantly, so the poses
 *               stay bit-identical with no broadcast.
 * No host synchronisation and no launch from Python inside the loop.
 * peer (NULL or world <= 1: single GPU, same result as frodo_gauss_newton):
 *   peer_buffers  device table [world] of the ranks' peer-mapped exchange buffers (e.g. torch
 *                 symmetric memory's buffer_ptrs_dev), local_buffer this rank's own;
 *                 each buffer holds 2 * max_edges * 35 doubles (blocks of even / odd epochs)
 *                 followed by 2 x 128 bytes of arrival counters, all ZERO when first used;
 *   epoch0        number of iterations ever run on these buffers (identical on all ranks),
 *   arrivals0[p]  arrivals so far on the counter of parity p (E per epoch of that parity);
 *                 the caller advances both after the call: epoch0 += max_iter, arrivals0[p] +=
 *                 E x (epochs of parity p in this call).  Iterations skipped after convergence
 *                 still arrive, so these counts do not depend on the data.
 * A wait that sees no progress for ~4 s gives up (dead peer), counts a failure in info[1] and
 * lets the kernel finish on what it has.
 * done (device int32, must be zero on entry) and info (device int32[2]) as in
 * frodo_gn_solve_retract; workspace of frodo_gn_workspace_bytes(K, E, N) bytes (ALL E edges). */
typedef struct frodo_gn_peer {
  const void* peer_buffers;
  void* local_buffer;
  int world;
  int rank;
  int max_edges;
  uint32_t epoch0;
  uint32_t arrivals0[2];
} frodo_gn_peer;
size_t frodo_gn_peer_buffer_bytes(int max_edges);
int frodo_gauss_newton_sharded(const frodo_gn_params* params, float* Twc, const float* Xs,
                               const float* Cs, const float* Kmat, const int32_t* ii_row,
                               const int32_t* jj_row, int E, int edge_lo, int edge_hi,
                               const int64_t* idx_local, const uint8_t* valid_local,
                               const float* Q_local, int K, int N, int num_fix, int max_iter,
                               float delta_thresh, const frodo_gn_peer* peer,
                               const frodo_gn_odom_terms* odom, float* dx, int32_t* done,
                               int32_t* info, void* workspace, size_t workspace_bytes,
                               frodo_stream_t stream);

int frodo_gn_peer_allreduce(const void* peer_buffers, const void* peer_signal_pads,
                            size_t buffer_offset_bytes, int signal_slot0, int rank, int world,
                            uint32_t epoch, double* out, size_t n, const int32_t* done,
                            frodo_stream_t stream);

/* ---- device-resident keyframe / edge arena --------------------------------
 * The reference re-stacks every keyframe's pointmap from CPU shared memory and every edge tensor
 * on each solve (FactorGraph.get_poses_points / prep_two_way_edges, global_opt.py:186-224;
 * SharedKeyframes, frame.py:374-390).  Here keyframes and directed edges live in caller-owned
 * device arrays with fixed capacity (host side: frodo-slam_b200/graph.py); a keyframe is packed
 * once when it is written, and a solve addresses the arena through three small index arrays.
 *
 * frodo_gn_pack_slot: X [N,3], C [N] of ONE keyframe -> row `slot` of a packed region of
 *   frodo_gn_packed_bytes(K_arena, N) bytes (same layout as frodo_gn_pack_points for K_arena).
 * frodo_gauss_newton_arena: the loop of frodo_gauss_newton over arena rows.
 *   Twc [K_arena,8], Xs [K_arena,N,3], Cs [K_arena,N], XC_packed (may be NULL): the arena;
 *   ii_slot, jj_slot [E] i32: arena rows of every directed edge's keyframes (data addressing);
 *   ii_rank, jj_rank [E] i32: their rank among the sorted unique keyframe ids of the edges
 *     (get_unique_kf_idx, global_opt.py:180-184): rows of the linear system, the first num_fix
 *     ranks are fixed;  pose_slot [K_solve] i32: rank -> arena row (where dx is applied);
 *   idx_ii2jj [E,N] i64, valid_match [E,N] u8, Q [E,N] f32: the edge arena (first E rows).
 *   odom (may be NULL; rays only): the gauss_newton_rays_odom terms, rows = arena rows.
 *   dx [K_solve - num_fix, 7], info as frodo_gauss_newton.
 *   workspace: frodo_gn_arena_workspace_bytes(K_arena, E, N). */
int frodo_gn_pack_slot(const float* X, const float* C, int K_arena, int slot, int N, float* XC,
                       frodo_stream_t stream);
size_t frodo_gn_arena_workspace_bytes(int K_arena, int E, int N);
int frodo_gauss_newton_arena(const frodo_gn_params* params, float* Twc, const float* Xs,
                             const float* Cs, const float* Kmat, const float* XC_packed,
                             const int32_t* ii_slot, const int32_t* jj_slot,
                             const int32_t* ii_rank, const int32_t* jj_rank,
                             const int32_t* pose_slot, const int64_t* idx_ii2jj,
                             const uint8_t* valid_match, const float* Q, int K_arena, int K_solve,
                             int N, int E, int num_fix, int max_iter, float delta_thresh,
                             const frodo_gn_odom_terms* odom, float* dx, int32_t* info,
                             void* workspace, size_t workspace_bytes, frodo_stream_t stream);

/* ---- the tracker's frame-to-keyframe pose solve -----------------------
 * FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 / solve, mast3r_slam/tracker.py:397-517
 * (+ geometry.py:17-104, nonlinear_optimizer.py:4-33): up to max_iters Gauss-Newton steps on the
 * relative pose T_CkCf = T_WCk^-1 T_WCf with Huber weights, stopped on device by the reference's
 * test (relative cost decrease < rel_error or ||tau|| < delta_norm).  One persistent cooperative
 * kernel, no host synchronisation.
 *   mode FRODO_GN_RAYS : residual [ray(3), distance(1)] of T_CkCf * Xf against Xk (sigma_a =
 *                        sigma_ray, sigma_b = sigma_dist).
 *   mode FRODO_GN_CALIB: residual [u, v, log z] of the pinhole projection of T_CkCf * Xf against
 *                        meas_k [N,3] (sigma_a = sigma_pixel, sigma_b = sigma_depth); needs Kmat
 *                        [3,3] (device), valid_meas_k [N] u8, height / width / pixel_border / z_eps.
 *   Xf [*,3] f32: the frame's points; row idx_f2k[n] is matched to keyframe point n (idx_f2k
 *     [N] i64, NULL = identity, i.e. Xf already gathered as tracker.py:395 does).
 *   Xk [N,3] f32 (rays only), Qk [N] f32, valid [N] u8: FULL-size arrays; the reference compacts
 *     by `valid` (tracker.py:293-295) or multiplies the information by it (:471-472), here a
 *     point with valid == 0 has zero weight.
 *   T_WCf, T_WCk [8] f32 device (t, q_xyzw, s).  Outputs (device): T_WCf_new [8] = T_WCk * T_CkCf,
 *     T_CkCf [8], info int32[2] = {iterations executed, 1 if the 7x7 system was not positive
 *     definite (the reference's "Cholesky failed")}, costs f64 [max_iters] (may be NULL).
 *   workspace: frodo_track_workspace_bytes(N) bytes. */
typedef struct frodo_track_params {
  int mode;
  float sigma_a;
  float sigma_b;
  float huber; /* 1.345 */
  int max_iters;
  float rel_error;
  float delta_norm;
  int height; /* calib only */
  int width;
  int pixel_border;
  float z_eps;
  int raw_calib; /* calib only, non-zero: Xf and Xk are the RAW pointmaps [H*W,3]; the kernel does
                    the calibrated branch of get_points_poses itself (tracker.py:383-393:
                    constrain_points_to_ray + meas_k = (u, v, log z), valid_meas_k = z > z_eps)
                    and meas_k / valid_meas_k are ignored (may be NULL) */
} frodo_track_params;

size_t frodo_track_workspace_bytes(int N);
int frodo_track_pose(const frodo_track_params* params, const float* Xf, const int64_t* idx_f2k,
                     const float* Xk, const float* Qk, const uint8_t* valid, const float* meas_k,
                     const uint8_t* valid_meas_k, const float* Kmat, const float* T_WCf,
                     const float* T_WCk, int N, float* T_WCf_new, float* T_CkCf, int32_t* info,
                     double* costs, void* workspace, size_t workspace_bytes,
                     frodo_stream_t stream);

/* frodo_track_pose without a workspace argument and without any allocation.  NO COUNTERPART IN
 * THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The scratch of the solve (grid-barrier words, per-CTA partial sums, the overflow record
 * buffer) lives in a static device array inside the library, sized for
 * frodo_track_pose_static_max_points() = 196,608 points (one 512x384 frame), 6.9 MB held by
 * every GPU that loads the library until the process exits.  N beyond that limit is rejected
 * with FRODO_ERR_ARG (use frodo_track_pose with your own workspace).  The call launches the same
 * kernel and calls no cudaMalloc / cudaFree / allocating library routine; inputs and outputs
 * are the caller's, exactly as for frodo_track_pose, and the results are bit-identical to it.
 * Every call on a device shares the one array: what must start at zero is cleared by each call
 * (stream-ordered), so back-to-back calls on one stream are safe, but calls that OVERLAP on the
 * GPU (different streams) race on it -- serialise them or use frodo_track_pose. */
int frodo_track_pose_static_max_points(void);
int frodo_track_pose_static(const frodo_track_params* params, const float* Xf,
                            const int64_t* idx_f2k, const float* Xk, const float* Qk,
                            const uint8_t* valid, const float* meas_k,
                            const uint8_t* valid_meas_k, const float* Kmat, const float* T_WCf,
                            const float* T_WCk, int N, float* T_WCf_new, float* T_CkCf,
                            int32_t* info, double* costs, frodo_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FRODO_B200_H */
