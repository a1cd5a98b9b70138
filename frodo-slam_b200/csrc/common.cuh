// Shared helpers for libfrodo_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/frodo_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libfrodo_b200 is written for sm_100a (B200) only"
#endif

namespace frodo {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs (compile-time model; see num_sms())
constexpr int kMaxDevices = 64;

// One-time per-DEVICE setup guard.  Function attributes (dynamic shared memory size, cluster
// size) belong to the device's context, so a process that drives several GPUs must set them
// once on each; a second thread racing here repeats an idempotent call, nothing worse.
struct PerDeviceOnce {
  volatile unsigned char set[kMaxDevices] = {};
  // Returns the current device in *dev (-1 on error) and whether setup is still needed there.
  bool needed(int* dev) {
    int d = -1;
    if (cudaGetDevice(&d) != cudaSuccess) d = -1;
    *dev = d;
    return d < 0 || d >= kMaxDevices || !set[d];
  }
  void mark(int dev) {
    if (dev >= 0 && dev < kMaxDevices) set[dev] = 1;
  }
};

// SM count of the current device (cached per device; 148 on a full B200, fewer on a MIG slice).
int num_sms();

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define FRODO_CUDA_TRY(expr)                                            \
  do {                                                                  \
    cudaError_t _e = (expr);                                            \
    if (_e != cudaSuccess) return ::frodo::cuda_fail(_e, #expr);        \
  } while (0)

#define FRODO_ARG_CHECK(cond)                                           \
  do {                                                                  \
    if (!(cond)) {                                                      \
      ::frodo::set_error("invalid argument: %s", #cond);                \
      return FRODO_ERR_ARG;                                             \
    }                                                                   \
  } while (0)

// Fused tails of the matching pipeline (frodo_match, matching.py:66-92): when set, the kernel
// writes the NEXT stage's outputs from its epilogue instead of its own.
struct IterProjTail {   // matching.py:66-75: p1 = trunc(p), valid = converged & |X11[p1] - X21| < thresh
  const float* X11;     // [B,H,W,3]; NULL = no tail (write p_new, converged)
  const float* X21;     // [B,H*W,3]
  float dist_thresh;
  int64_t* p1;          // [B,H*W,2]
  uint8_t* valid;       // [B,H*W]
};
struct RefineTail {     // matching.py:87-92: valid &= score > threshold (fp16 compare), idx = x + W y
  const uint8_t* valid_in;  // NULL = no tail (write p1_new, scores)
  float threshold;
  int64_t* idx;         // [B,N]
  uint8_t* valid;       // [B,N]
};

// Launchers implemented in the .cu files (host side, return FRODO_* codes).
int launch_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_new,
                     uint8_t* converged, int B, int H, int W, int N, int max_iter,
                     float lambda_init, float cost_thresh, cudaStream_t s,
                     const IterProjTail* tail = nullptr);
int launch_refine_f16(const __half* D11, const __half* D21, const int64_t* p1, int64_t* p1_new,
                      __half* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s, const float* D21_f32 = nullptr,
                      const RefineTail* tail = nullptr);
int refine_forced_patch(int set);
bool refine_f16_takes_tile_kernel(const void* D11, const void* D21, const void* p1,
                                  const void* p1_new, int B, int H, int W, int N, int F);
int launch_refine_f32(const float* D11, const float* D21, const int64_t* p1, int64_t* p1_new,
                      float* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s);

int launch_refine_f64(const double* D11, const double* D21, const int64_t* p1, int64_t* p1_new,
                      double* scores, int B, int H, int W, int N, int F, int radius, int dil,
                      cudaStream_t s);

int launch_match_prep(const float* X11, const float* X21, const int64_t* idx_init,
                      float* rays_with_grad, float* pts3d_norm, float* p_init, int B, int H, int W,
                      cudaStream_t s);
int launch_match_occlusion(const float* X11, const float* X21, const float* p_iter,
                           const uint8_t* conv, float dist_thresh, int64_t* p1, uint8_t* valid,
                           int B, int H, int W, cudaStream_t s);
int launch_convert_f32_f16(const float* src, __half* dst, size_t n, cudaStream_t s);
int launch_constrain_points(const float* X, const float* Kmat, int num_frames, int H, int W,
                            float* out, cudaStream_t s);
int launch_match_finish(const int64_t* p1, const void* scores, bool scores_f16,
                        const uint8_t* valid_in, float threshold, int64_t* idx, uint8_t* valid,
                        int B, int N, int W, cudaStream_t s);


int launch_match_quality(const float* Q_src, const float* Q_dst, const float* C_src,
                         const float* C_dst, const int64_t* idx, const uint8_t* valid_match,
                         float Q_conf, float C_conf, float* Q_out, uint8_t* valid_opt,
                         uint8_t* valid_q, int32_t* counts, int B, int N, int N_src,
                         cudaStream_t s);

int launch_peer_allreduce(const void* bufs_dev, const void* pads_dev, size_t buf_offset,
                          int pad_slot0, int rank, int world, uint32_t epoch, double* out, size_t n,
                          const int32_t* done, cudaStream_t s);

size_t track_workspace_bytes(int N);
int launch_track_pose(const frodo_track_params& prm, const float* Xf, const int64_t* idx,
                      const float* Xk, const float* Qk, const uint8_t* valid, const float* meas,
                      const uint8_t* valid_meas, const float* Kmat, const float* T_WCf,
                      const float* T_WCk, int N, float* T_WCf_out, float* T_CkCf_out,
                      int32_t* info, double* costs, void* workspace, cudaStream_t s);

int track_static_max_points();
int launch_track_pose_static(const frodo_track_params& prm, const float* Xf, const int64_t* idx,
                             const float* Xk, const float* Qk, const uint8_t* valid,
                             const float* meas, const uint8_t* valid_meas, const float* Kmat,
                             const float* T_WCf, const float* T_WCk, int N, float* T_WCf_out,
                             float* T_CkCf_out, int32_t* info, double* costs, cudaStream_t s);

}  // namespace frodo
