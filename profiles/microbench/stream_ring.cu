// Read-only streaming ceiling on B200: which way of reading a >L2 buffer reaches the HBM peak?
//   mode 0: cp.async.bulk ring (stages x stage_bytes per CTA, one issuing thread, full/empty
//           mbarriers), consumers touch 16 B per thread per stage
//   mode 1: plain LDG.128, UNROLL independent loads per thread, grid-stride
//   mode 2: like 0 but each stage is split into `pieces` separate bulk copies from `pieces`
//           separate arrays (the shape of the GN accumulate kernel: idx/XP/Q/valid)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../frodo-slam_b200/csrc
//        stream_ring.cu -o stream_ring
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "ptx_async.cuh"
using namespace frodo::ptx;

// piece sizes as fractions of the stage: equal, or the GN tile's idx/XP/Q/valid = 8:16:4:1 of 29
__global__ void ring_kernel(const char* __restrict__ src, size_t total, int stage_bytes, int stages,
                            int pieces, unsigned* sink, int multi_issue = 0, int gn_shape = 0) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + 16;
  unsigned char* buf = smem + 256;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full + s, multi_issue ? pieces : 1);
      mbar_init(empty + s, blockDim.x);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const size_t ntiles = total / stage_bytes;
  const size_t per = (ntiles + gridDim.x - 1) / gridDim.x;
  const size_t t0 = per * blockIdx.x, t1 = min(ntiles, t0 + per);
  int pb[4], po[4];   // piece bytes / offsets inside a stage
  if (gn_shape) {
    const int u = stage_bytes / 29;
    pb[0] = 8 * u; pb[1] = 16 * u; pb[2] = 4 * u; pb[3] = u;
  } else {
    for (int p = 0; p < 4; ++p) pb[p] = stage_bytes / pieces;
  }
  po[0] = 0;
  for (int p = 1; p < 4; ++p) po[p] = po[p - 1] + pb[p - 1];
  const size_t piece_stride = (total / 4) & ~(size_t)127;   // pieces come from separate arrays
  auto issue_piece = [&](size_t t, int s, int p) {
    bulk_g2s(buf + (size_t)s * stage_bytes + po[p], src + p * piece_stride + t * pb[p], pb[p], full + s);
  };
  auto issue = [&](size_t t, int s) {
    mbar_arrive_expect_tx(full + s, stage_bytes);
    for (int p = 0; p < pieces; ++p) issue_piece(t, s, p);
  };
  const int my_piece = (tid & 31) == 0 ? tid >> 5 : -1;   // lane 0 of warp p issues piece p
  if (multi_issue) {
    if (my_piece >= 0 && my_piece < pieces)
      for (int s = 0; s < stages && t0 + s < t1; ++s) {
        mbar_arrive_expect_tx(full + s, pb[my_piece]);
        issue_piece(t0 + s, s, my_piece);
      }
  } else if (tid == 0)
    for (int s = 0; s < stages && t0 + s < t1; ++s) issue(t0 + s, s);
  unsigned acc = 0;
  int s = 0; unsigned ph = 0;
  for (size_t t = t0; t < t1; ++t) {
    mbar_wait(full + s, ph);
    acc += *reinterpret_cast<const unsigned*>(buf + (size_t)s * stage_bytes + (tid * 16) % stage_bytes);
    mbar_arrive(empty + s);
    if (multi_issue) {
      if (my_piece >= 0 && my_piece < pieces && t + stages < t1) {
        mbar_wait(empty + s, ph);
        mbar_arrive_expect_tx(full + s, pb[my_piece]);
        issue_piece(t + stages, s, my_piece);
      }
    } else if (tid == 0 && t + stages < t1) {
      mbar_wait(empty + s, ph);
      issue(t + stages, s);
    }
    if (++s == stages) { s = 0; ph ^= 1; }
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <int UNROLL>
__global__ void ldg_kernel(const uint4* __restrict__ src, size_t n16, unsigned* sink) {
  unsigned acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (UNROLL - 1) * stride < n16; i += UNROLL * stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) v[u] = __ldcs(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

int main(int argc, char** argv) {
  const size_t total = (size_t)2304 << 20;   // 2.25 GiB, the GN iteration's footprint
  char* src; unsigned* sink;
  cudaMalloc(&src, total); cudaMalloc(&sink, 4);
  cudaMemset(src, 1, total);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto time = [&](auto&& launch, const char* name) {
    for (int i = 0; i < 2; ++i) launch();
    cudaEventRecord(e0);
    const int reps = 5;
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    printf("%-58s %8.1f us  %7.1f GB/s %s\n", name, ms / reps * 1e3, total / (ms / reps * 1e-3) / 1e9,
           err == cudaSuccess ? "" : cudaGetErrorString(err));
  };
  cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  char name[128];
  const int cfgs[][7] = {
      // stage_bytes, stages, pieces, ctas/SM, threads, multi_issue, gn_shape
      {14848, 5, 1, 1, 256, 0, 0}, {14848, 5, 4, 1, 256, 0, 0}, {14848, 5, 4, 1, 256, 1, 0},
      {14848, 5, 4, 1, 256, 0, 1}, {14848, 5, 4, 1, 256, 1, 1}, {14848, 8, 4, 1, 256, 0, 1},
      {14848, 12, 4, 1, 256, 0, 1}, {14848, 12, 4, 1, 256, 1, 1},
      {29696, 4, 4, 1, 256, 0, 1}, {29696, 4, 4, 1, 256, 1, 1}, {29696, 6, 4, 1, 256, 0, 1},
      {29696, 6, 4, 1, 256, 1, 1}, {29696, 3, 4, 1, 256, 0, 1},
      {59392, 3, 4, 1, 256, 0, 1}, {59392, 3, 4, 1, 256, 1, 1},
      {14848, 5, 4, 2, 256, 0, 1}, {14848, 5, 4, 2, 256, 1, 1}, {14848, 3, 4, 2, 256, 0, 1},
      {29696, 3, 4, 2, 256, 0, 1}, {7424, 6, 4, 4, 128, 0, 1}, {7424, 6, 4, 4, 128, 1, 1},
      {14848, 3, 4, 4, 128, 0, 1}, {14848, 3, 4, 4, 128, 1, 1},
  };
  for (auto& c : cfgs) {
    const int sb = c[0], st = c[1], pc = c[2], cps = c[3], th = c[4], mi = c[5], gs = c[6];
    const size_t sm = 256 + (size_t)sb * st;
    snprintf(name, sizeof name, "ring %5d x%2d pc %d cta/SM %d (%3zu KB/SM) %s %s", sb, st, pc,
             cps, sm * cps >> 10, mi ? "4-issuer" : "1-issuer", gs ? "8:16:4:1" : "equal");
    time([&] { ring_kernel<<<148 * cps, th, sm>>>(src, total, sb, st, pc, sink, mi, gs); }, name);
  }
  const size_t n16 = total / 16;
  time([&] { ldg_kernel<4><<<148 * 8, 256>>>((const uint4*)src, n16, sink); }, "LDG.128 x4, 8 CTAs/SM x 256");
  time([&] { ldg_kernel<8><<<148 * 8, 256>>>((const uint4*)src, n16, sink); }, "LDG.128 x8, 8 CTAs/SM x 256");
  time([&] { ldg_kernel<8><<<148 * 4, 512>>>((const uint4*)src, n16, sink); }, "LDG.128 x8, 4 CTAs/SM x 512");
  time([&] { ldg_kernel<16><<<148 * 4, 256>>>((const uint4*)src, n16, sink); }, "LDG.128 x16, 4 CTAs/SM x 256");
  time([&] { cudaMemsetAsync(src, 2, total); }, "cudaMemset (write only)");
  return 0;
}
