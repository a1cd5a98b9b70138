// Dependent-issue latency of fp64 operations on sm_100a (one warp, one chain).
#include <cstdio>
#include <cuda_runtime.h>
#define N 2048
template <int V>
__global__ void k(double* out, double seed, long long* cyc) {
  double x = seed + threadIdx.x * 1e-3, y = seed * 0.5, c = 1.0000001;
  float f = (float)seed;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) {
    if (V == 0) x = fma(x, c, y);
    if (V == 1) x = x * c;
    if (V == 2) x = x + y;
    if (V == 3) asm volatile("rsqrt.approx.ftz.f64 %0, %0;" : "+d"(x));
    if (V == 4) {  // rsqrt_pos of gn_chol_cluster.cu
      double r;
      asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
      const double e = fma(x, -(r * r), 1.0);
      const double p = fma(e, 0.375, 0.5);
      x = fma(p, r * e, r) + 1.0;
    }
    if (V == 5) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(1.0001f), "f"(0.5f));
    if (V == 6) {  // fp32 seed + one fp64 Newton step
      float s;
      asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"((float)x));
      const double r = (double)s;
      const double e = fma(x, -(r * r), 1.0);
      x = fma(0.5 * r, e, r) + 1.0;
    }
    if (V == 7) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 1) & 31);
    if (V == 8) asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x));
  }
  long long t1 = clock64();
  out[threadIdx.x] = x + f;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}
template <int V> void run(const char* name) {
  double* out; long long* cyc; long long h;
  cudaMalloc(&out, 256); cudaMalloc(&cyc, 8);
  k<V><<<1, 32>>>(out, 1.5, cyc); k<V><<<1, 32>>>(out, 1.5, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-44s %7.1f cycles per dependent op\n", name, (double)h / N);
}
int main() {
  run<0>("DFMA"); run<1>("DMUL"); run<2>("DADD"); run<3>("MUFU.RSQ64H (rsqrt.approx.f64)");
  run<4>("rsqrt_pos (seed + cubic step) + DADD"); run<5>("FFMA");
  run<6>("fp32 rsqrt seed + 1 fp64 Newton + DADD"); run<7>("SHFL of a double");
  run<8>("rcp.approx.f64");
  return 0;
}
