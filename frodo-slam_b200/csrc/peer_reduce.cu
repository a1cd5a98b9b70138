// All-reduce of the assembled Gauss-Newton system [H | g] (fp64, 1.56 MB for 64 keyframes) over
// NVLink peer memory, for the edge-sharded multi-GPU driver (frodo-slam_b200/dist.py).
//
// The exchange is small and latency-bound, so it is ONE kernel per iteration instead of a NCCL
// collective: every rank keeps its partial system in a symmetric (peer-mapped) buffer, and
//   1. block 0 publishes "my partial of this epoch is complete" into every peer's signal pad
//      (system-scope release), every block waits until all peers have published (acquire);
//   2. each thread sums its elements over the ranks IN RANK ORDER 0..W-1 with 16-byte loads
//      straight from the peers' memory and writes the total to a local buffer.
// The rank order is the same on every GPU, so all ranks end up with bit-identical systems and the
// replicated Cholesky keeps the poses bit-identical with no broadcast (NCCL gives the same
// guarantee; an order that started at the local rank would not).  The symmetric buffers are
// double-buffered by epoch parity, which makes one barrier per iteration sufficient: a rank can
// only overwrite the buffer of epoch e when it assembles epoch e+2, i.e. after it completed the
// reduction of epoch e+1, which every peer entered only after finishing its reads of epoch e.
// `done` (the solver's convergence flag, identical on all ranks) skips the whole step.
#include "gn_internal.cuh"

namespace frodo {

namespace {

constexpr int kPeerThreads = 256;
constexpr int kMaxWorld = 16;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kPeerThreads)
peer_allreduce_kernel(const double* const* __restrict__ bufs, uint32_t* const* __restrict__ pads,
                      size_t buf_offset, int pad_slot0, int rank, int world, uint32_t epoch,
                      double* __restrict__ out, size_t n, const int32_t* __restrict__ done) {
  if (done != nullptr && *done != 0) return;
  const int tid = threadIdx.x;
  // ---- cross-rank barrier on the signal pads: slot [pad_slot0 + r] of rank q's pad = the
  // latest epoch rank r has published to q
  if (blockIdx.x == 0 && tid < world) {
    __threadfence_system();
    st_release_sys(pads[tid] + pad_slot0 + rank, epoch);
  }
  if (tid < world) {
    const uint32_t* mine = pads[rank] + pad_slot0 + tid;
    const long long t0 = clock64();  // bounded: a dead peer must not hang this GPU
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0 && clock64() - t0 < 8000000000LL) {
    }
  }
  __syncthreads();
  // ---- fixed-order sum over the ranks, 16 bytes per load
  const double* src[kMaxWorld];
#pragma unroll
  for (int r = 0; r < kMaxWorld; ++r)
    src[r] = r < world ? reinterpret_cast<const double*>(
                             reinterpret_cast<const char*>(bufs[r]) + buf_offset)
                       : nullptr;
  const size_t n2 = n / 2;
  const size_t stride = (size_t)gridDim.x * kPeerThreads;
  for (size_t e = (size_t)blockIdx.x * kPeerThreads + tid; e < n2; e += stride) {
    double2 v[kMaxWorld];
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      if (r < world) v[r] = __ldcg(reinterpret_cast<const double2*>(src[r]) + e);
    double2 acc = v[0];
#pragma unroll
    for (int r = 1; r < kMaxWorld; ++r)
      if (r < world) {
        acc.x += v[r].x;
        acc.y += v[r].y;
      }
    reinterpret_cast<double2*>(out)[e] = acc;
  }
  if ((n & 1) && blockIdx.x == 0 && tid == 0) {
    double acc = src[0][n - 1];
#pragma unroll
    for (int r = 1; r < kMaxWorld; ++r)
      if (r < world) acc += src[r][n - 1];
    out[n - 1] = acc;
  }
}

// Consumer-side wait of the fused exchange (gn_accumulate.cu pushes per-edge blocks to every
// rank and bumps its arrival counter): used by the solve paths that do not wait inside their own
// kernel (small systems, the single-CTA fallback).  Bounded spin: a dead peer must not hang the GPU.
__global__ void peer_wait_kernel(const uint32_t* __restrict__ counter, uint32_t target,
                                 const int32_t* __restrict__ done, int32_t* __restrict__ info) {
  if (done != nullptr && *done != 0) return;
  const long long t0 = clock64();
  uint32_t v;
  for (;;) {
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    if ((int32_t)(v - target) >= 0) break;
    if (clock64() - t0 > 8000000000LL) {  // ~4 s
      if (info != nullptr) atomicAdd(info + 1, 1);
      break;
    }
    __nanosleep(40);
  }
  v = ld_acquire_sys(counter);
}

}  // namespace

int launch_gn_peer_wait(const uint32_t* counter, uint32_t target, const int32_t* done,
                        int32_t* info, cudaStream_t s) {
  peer_wait_kernel<<<1, 1, 0, s>>>(counter, target, done, info);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

int launch_peer_allreduce(const void* bufs_dev, const void* pads_dev, size_t buf_offset,
                          int pad_slot0, int rank, int world, uint32_t epoch, double* out, size_t n,
                          const int32_t* done, cudaStream_t s) {
  if (n == 0) return FRODO_OK;
  int dev = 0, sms = 0;
  FRODO_CUDA_TRY(cudaGetDevice(&dev));
  FRODO_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // one wave: every block spins on the signal pad, so all of them must be resident
  size_t want = (n / 2 + kPeerThreads - 1) / kPeerThreads;
  if (want < 1) want = 1;
  const int blocks = (int)(want < (size_t)sms ? want : (size_t)sms);
  peer_allreduce_kernel<<<blocks, kPeerThreads, 0, s>>>(
      static_cast<const double* const*>(bufs_dev), static_cast<uint32_t* const*>(pads_dev),
      buf_offset, pad_slot0, rank, world, epoch, out, n, done);
  FRODO_CUDA_TRY(cudaGetLastError());
  return FRODO_OK;
}

}  // namespace frodo
