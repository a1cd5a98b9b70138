// Internal declarations shared by the Gauss-Newton translation units.
#pragma once
#include "common.cuh"

namespace frodo {

constexpr int kGnBlock = 35;  // per-edge block: 28 upper-tri entries of A + 7 of g_j

// Points handled by one accumulate CTA for a graph of E edges x N points.
int gn_pts_per_cta(int E, int N);

struct GnWorkspace {
  int32_t* ii_row;     // [E]
  int32_t* jj_row;     // [E]
  int32_t* scalars;    // [8]: 0 n_unique, 1 done, 2..3 info
  int32_t* kf_need;    // [K]: launch_gn_mark_keyframes (edge-sharded solves)
  int32_t* tickets;    // [E]  (zero between launches)
  float* partials;     // [E * nchunks * 35]
  double* edge_blocks; // [E * 35]
  double* Hg;          // [(n+1) * n]
  double* chol;        // scratch of the dataflow Cholesky (gn_chol_scratch_bytes)
  float* packed;       // [2 * K * N * 4]: XC float4 (x,y,z,c) then XP pair-interleaved
  size_t packed_off;   // byte offset of `packed` inside the workspace
  size_t bytes;
};
// Carves `base` (may be NULL to only compute sizes).  Returns total bytes.
size_t gn_carve_workspace(void* base, int K, int E, int N, GnWorkspace* ws);

// q_ids [nq] (optional): extra ids ranked against the same sorted unique list -> q_rows.
int launch_gn_edge_index(const int64_t* ii, const int64_t* jj, int E, int32_t* ii_row,
                         int32_t* jj_row, int32_t* n_unique, cudaStream_t s,
                         const int64_t* q_ids = nullptr, int nq = 0, int32_t* q_rows = nullptr);

// gauss_newton_rays_odom extras (gn_odom.cu).  Scratch: [2*Eo] int32 rows (odom_ii ranks then
// odom_jj ranks, written by launch_gn_edge_index) followed, 256-byte aligned, by [Eo][210] fp64.
struct GnOdom {
  const int64_t* odom_ii;  // [Eo] keyframe ids, used raw as Twc rows (gn_kernels.cu:1121-1122)
  const int64_t* odom_jj;
  const float* Tij;        // [Eo, stride]: t(3), q_xyzw(4)
  int stride;
  int Eo;
  const double* s_bar;     // [K] device, may be NULL
  float sigma_t, sigma_r, sigma_scale;
};
size_t gn_odom_scratch_bytes(int Eo);
size_t gn_odom_blocks_bytes(int Eo);
int launch_gn_add_odom_rows(const GnOdom& od, const float* Twc, int K, int num_fix, double* Hg,
                            const int32_t* done, const int32_t* rows, double* blocks,
                            const int32_t* pose_rows, cudaStream_t s);
int launch_gn_add_odom(const GnOdom& od, const float* Twc, int K, int num_fix, double* Hg,
                       const int32_t* done, void* scratch, cudaStream_t s);
// Bytes of the packed pointmap region (XC [K][N] float4 + XP padded to 512-point blocks).
size_t gn_packed_bytes(int K, int N);
// `need` (may be null): per keyframe, bit 0 = its XC row is read (it is the i of an edge of this
// rank), bit 1 = its XP row is (it is the j); rows nobody reads are not packed.
int launch_gn_pack_points(const float* Xs, const float* Cs, int K, int N, float* XC,
                          cudaStream_t s, const int32_t* need = nullptr);
// need[k] for the keyframe rows 0..K-1 from the edges [0, E) of ii_row / jj_row.
int launch_gn_mark_keyframes(const int32_t* ii_row, const int32_t* jj_row, int E, int K,
                             int32_t* need, cudaStream_t s);
// One keyframe (X [N,3], C [N]) into row `slot` of a packed region laid out for K_total keyframes.
int launch_gn_pack_slot(const float* X, const float* C, int K_total, int slot, int N, float* XC,
                        cudaStream_t s);
// Multi-GPU output of the accumulate kernels (world = 0: plain local edge_blocks).  bufs[r]
// = base of rank r's peer-mapped exchange buffer; the per-edge blocks of the current epoch live
// at eb_offset inside it, indexed by GLOBAL edge number (edge_base + local edge), the arrival
// counter (u32) at cnt_offset.
struct GnPeerOut {
  void* const* bufs;  // device table [world]
  size_t eb_offset;
  size_t cnt_offset;
  int world;
  int edge_base;
};
// XC: pointmaps packed by launch_gn_pack_points (XC [K][N] float4, then XP), or NULL.
int launch_gn_accumulate(const frodo_gn_params& prm, const float* Twc, const float* Xs,
                         const float* Cs, const float* Kmat, const float* XC,
                         const int32_t* ii_row,
                         const int32_t* jj_row, const int64_t* idx, const uint8_t* valid,
                         const float* Q, int K, int N, int E, float* partials, int32_t* tickets,
                         double* edge_blocks, const int32_t* done, cudaStream_t s,
                         const GnPeerOut* peer_out = nullptr);
int launch_gn_assemble(const double* edge_blocks, const int32_t* ii_row, const int32_t* jj_row,
                       int E, int K, int num_fix, double* Hg, const int32_t* done, cudaStream_t s);
// Where the normal equations of one iteration are: the dense fp64 [H | g] (`Hg`, valid when
// `dense` is set) or the per-edge blocks the accumulate kernel left (`edge_blocks` + the edge
// ranks).  `Hg` must always point at an [(n+1) x n] buffer: paths that need the dense form build
// it there.
struct GnSystem {
  double* Hg;
  bool dense;
  const double* edge_blocks;
  const int32_t* ii_row;  // negative rows mark padding entries (skipped)
  const int32_t* jj_row;
  int E;
  // multi-GPU: the blocks are complete once *wait_counter has reached wait_target (arrivals of
  // every rank's accumulate kernel, gn_accumulate.cu); NULL = no wait
  const uint32_t* wait_counter = nullptr;
  uint32_t wait_target = 0;
};
// n_max: the system size the Cholesky scratch was laid out and initialised for (7 x keyframes).
int launch_gn_solve_retract(const GnSystem& sys, float* Twc, float* dx, int K, int num_fix,
                            float delta_thresh, int32_t* done, int32_t* info, double* chol_scratch,
                            int n_max, cudaStream_t s, const int32_t* pose_rows = nullptr);
size_t gn_chol_scratch_bytes(int n);
size_t gn_chol_stamps_offset(int n);
int gn_chol_scratch_init(double* scratch, int n_max, cudaStream_t s);
int launch_chol_flow(const double* Hg, const double* edge_blocks, const int32_t* ii_row,
                     const int32_t* jj_row, int E, float* Twc, float* dx, int K, int num_fix,
                     float delta_thresh, int32_t* done, int32_t* info, double* chol_scratch,
                     int n_max, int* handled, cudaStream_t s, const int32_t* pose_rows = nullptr,
                     const uint32_t* wait_counter = nullptr, uint32_t wait_target = 0);
// Spins (one thread, bounded) until *counter has reached target; on time-out bumps info[1].
int launch_chol_small(const double* edge_blocks, const int32_t* ii_row, const int32_t* jj_row,
                      int E, float* Twc, float* dx, int K, int num_fix, float delta_thresh,
                      int32_t* done, int32_t* info, int* handled, cudaStream_t s,
                      const int32_t* pose_rows);
int launch_gn_peer_wait(const uint32_t* counter, uint32_t target, const int32_t* done,
                        int32_t* info, cudaStream_t s);

}  // namespace frodo
