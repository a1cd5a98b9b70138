// Microbenchmark of the diagonal-tile factor + inversion of gn_chol_flow.cu in isolation:
// (a) one warp alone, (b) one warp while 7 others wait at a barrier, (c) the same with the
// function's global stores going to a device buffer.  Prints clock64 cycles per call.
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../frodo-slam_b200/csrc \
//      diag_factor.cu -o diag_factor
#include <cstdio>
#include <vector>
#include <cmath>
#define FRODO_MICROBENCH 1
namespace frodo { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*) { return 3; }
int num_sms() { return 148; } }
#include "gn_chol_flow.cu"

using namespace frodo;

__global__ void __launch_bounds__(256, 1) bench(const double* tile_g, double* Wg, double* ytail, long long* out, int reps,
                      int mode, const uint32_t* flag) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  double* D = sm, *cols = sm + 1024, *Wl = sm + 2048;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[3072 + i] = 1e-3 * i;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) D[i] = tile_g[i];
  __syncthreads();
  for (int rep = 0; rep < reps; ++rep) {
    long long t0 = clock64();
    if (threadIdx.x < 32) {
      bool ok = diag_factor_invert_warp(D, cols, Wl, Wg, 32, -1, ytail);
      if (threadIdx.x == 0) { out[rep] = clock64() - t0; out[64 + rep] = ok; }
    } else if (mode >= 3) {
      // fp64 load on other warps while warp 0 factors: mode 3 = warps 2,3,6,7 (sub-partitions
      // 2,3), mode 4 = warps 4 (same sub-partition as warp 0), mode 5 = warps 1..7
      const int w = threadIdx.x >> 5;
      const bool on = (mode == 3) ? (w == 2 || w == 3 || w == 6 || w == 7) : (mode == 4) ? (w == 4) : true;
      if (on) {
        double acc[4][4] = {};
        const int tg = threadIdx.x & 63;
        for (int it = 0; it < 6; ++it) tile_mma(sm + 3072, sm + 3072, tg >> 3, tg & 7, 0, 32, acc);
        double sacc = 0; for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) sacc += acc[a][b];
        if (sacc == 1234.5) out[100] = 1;
        if ((threadIdx.x & 31) == 0) out[80 + w] = clock64() - t0;
      }
    } else if (mode == 1 && threadIdx.x == 32) {
      for (int it = 0; it < 40; ++it) { if (ld_relaxed(flag) == 12345u) break; __nanosleep(32); }
    } else if (mode == 2 && threadIdx.x == 32) {
      for (int it = 0; it < 200; ++it) { uint32_t v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); if (v == 12345u) break; }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256, 1) pollers(const uint32_t* flag, int mode) {
  extern __shared__ unsigned char dummy[];
  if (threadIdx.x == 0) {
    if (mode == 0) { while (ld_relaxed(flag) != 777u) __nanosleep(32); }
    else { while (ld_relaxed(flag) != 777u) { } }
  }
  __syncthreads();
}

int main() {
  std::vector<double> A(1024), S(1024);
  // SPD tile: A = M M^T + 32 I, stored swizzled column-major
  std::vector<double> M(1024);
  for (int i = 0; i < 1024; ++i) M[i] = std::sin(0.37 * i) ;
  for (int r = 0; r < 32; ++r)
    for (int c = 0; c < 32; ++c) {
      double s = (r == c) ? 32.0 : 0.0;
      for (int k = 0; k < 32; ++k) s += M[r * 32 + k] * M[c * 32 + k];
      S[swz(r, c)] = s;
    }
  double *d_tile, *d_W, *d_y; long long* d_out;
  cudaMalloc(&d_tile, 8192); cudaMalloc(&d_W, 8192); cudaMalloc(&d_y, 256); cudaMalloc(&d_out, 128 * 8);
  cudaMemcpy(d_tile, S.data(), 8192, cudaMemcpyHostToDevice);
  uint32_t* d_flag; cudaMalloc(&d_flag, 256); cudaMemset(d_flag, 0, 256);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  struct V { int nt, mode; size_t smem; const char* name; } vs[] = {
    {32, 0, 4 * 8192, "1 warp, 32 KB smem"}, {256, 0, 4 * 8192, "8 warps (7 parked), 32 KB"},
    {256, 0, 220 * 1024, "8 warps (7 parked), 220 KB smem"},
    {256, 1, 220 * 1024, "warp 1 polls relaxed+nanosleep, 220 KB"},
    {256, 2, 220 * 1024, "warp 1 polls ld.acquire (CCTL.IVALL), 220 KB"},
    {256, 3, 220 * 1024, "warps 2,3,6,7 run fp64 tile products"},
    {256, 4, 220 * 1024, "warp 4 (same sub-partition) runs fp64 tile products"},
    {256, 5, 220 * 1024, "warps 1..7 run fp64 tile products"}};
  cudaFuncSetAttribute(pollers, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  cudaStream_t s1, s2; cudaStreamCreate(&s1); cudaStreamCreate(&s2);
  for (int pm = 0; pm < 2; ++pm) {
    cudaMemset(d_flag, 0, 4);
    pollers<<<140, 256, 220 * 1024, s1>>>(d_flag, pm);
    bench<<<1, 256, 220 * 1024, s2>>>(d_tile, d_W, d_y, d_out, 6, 0, d_flag + 8);
    cudaStreamSynchronize(s2);
    uint32_t go = 777u; cudaMemcpyAsync(d_flag, &go, 4, cudaMemcpyHostToDevice, s2);
    cudaError_t e = cudaDeviceSynchronize();
    long long out[128];
    cudaMemcpy(out, d_out, sizeof out, cudaMemcpyDeviceToHost);
    printf("%-48s (%s): cycles per call:", pm ? "140 SMs polling one flag (tight)" : "140 SMs polling one flag (nanosleep 32)", cudaGetErrorString(e));
    for (int i = 0; i < 6; ++i) printf(" %lld", out[i]);
    printf("\n");
  }
  for (auto& v : vs) {
    const int nt = v.nt;
    for (int z = 0; z < 128; ++z) {} bench<<<1, nt, v.smem>>>(d_tile, d_W, d_y, d_out, 6, v.mode, d_flag);
    cudaError_t e = cudaDeviceSynchronize();
    long long out[128];
    cudaMemcpy(out, d_out, sizeof out, cudaMemcpyDeviceToHost);
    printf("%-48s (%s): cycles per call:", v.name, cudaGetErrorString(e));
    for (int i = 0; i < 6; ++i) printf(" %lld", out[i]);
    printf("  ok=%lld", out[64]);
    if (v.mode >= 3) { printf("  other warps' cycles:"); for (int w = 1; w < 8; ++w) printf(" %lld", out[80 + w]); }
    printf("\n");
  }
  return 0;
}
