// Microbenchmark of the diagonal-tile factor + inversion of gn_chol_flow.cu in isolation
// (four warps, one per SM sub-partition), with and without the other four warps of the CTA
// doing fp64 tile products next to it.  Prints clock64 cycles per call.
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
//      -I../../frodo-slam_b200/csrc diag_factor.cu -o diag_factor
#include <cmath>
#include <cstdio>
#include <vector>
namespace frodo { void set_error(const char*, ...) {} int cuda_fail(cudaError_t, const char*) { return 3; }
int num_sms() { return 148; } }
#include "gn_chol_flow.cu"

using namespace frodo;

__global__ void __launch_bounds__(256, 1) bench(const double* tile_g, double* ytail, long long* out,
                                                int reps, int mode) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sm = reinterpret_cast<double*>(smem_raw);
  double *D = sm, *cols = sm + 1024, *Wl = sm + 2048;
  __shared__ int s_ok;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
    sm[3072 + i] = 1e-3 * i;
    D[i] = tile_g[i];
  }
  __syncthreads();
  for (int rep = 0; rep < reps; ++rep) {
    const long long t0 = clock64();
    if (threadIdx.x < 128) {
      const bool ok = diag_factor_invert(D, cols, Wl, 32, -1, ytail, &s_ok);
      if (threadIdx.x == 0) { out[rep] = clock64() - t0; out[64 + rep] = ok; }
    } else if (mode == 1) {  // warps 4-7 run fp64 tile products meanwhile
      double acc[4][4] = {};
      const int tg = threadIdx.x & 63;
      for (int it = 0; it < 4; ++it) tile_mma(sm + 3072, sm + 3072, tg >> 3, tg & 7, 0, 32, acc);
      double sacc = 0;
      for (int a = 0; a < 4; ++a) for (int b = 0; b < 4; ++b) sacc += acc[a][b];
      if (sacc == 1234.5) out[100] = 1;
    }
    __syncthreads();
  }
}

int main() {
  std::vector<double> S(1024), M(1024);
  for (int i = 0; i < 1024; ++i) M[i] = std::sin(0.37 * i);
  for (int r = 0; r < 32; ++r)
    for (int c = 0; c < 32; ++c) {  // SPD tile M M^T + 32 I, swizzled column-major
      double s = (r == c) ? 32.0 : 0.0;
      for (int k = 0; k < 32; ++k) s += M[r * 32 + k] * M[c * 32 + k];
      S[swz(r, c)] = s;
    }
  double *d_tile, *d_y;
  long long* d_out;
  cudaMalloc(&d_tile, 8192); cudaMalloc(&d_y, 256); cudaMalloc(&d_out, 128 * 8);
  cudaMemcpy(d_tile, S.data(), 8192, cudaMemcpyHostToDevice);
  const char* names[] = {"4 factor warps, 4 parked", "4 factor warps, 4 running fp64 tile products"};
  for (int mode = 0; mode < 2; ++mode) {
    bench<<<1, 256, 4 * 8192>>>(d_tile, d_y, d_out, 6, mode);
    cudaError_t e = cudaDeviceSynchronize();
    long long out[128];
    cudaMemcpy(out, d_out, sizeof out, cudaMemcpyDeviceToHost);
    printf("%-48s (%s): cycles per call:", names[mode], cudaGetErrorString(e));
    for (int i = 0; i < 6; ++i) printf(" %lld", out[i]);
    printf("  ok=%lld\n", out[64]);
  }
  return 0;
}
