// Which issue pipes do the packed-half instructions of refine_matches use on sm_100a?
// Each variant runs 8 independent dependency chains per thread; cycles per warp-instruction
// per SM sub-partition are printed.   nvcc -arch=sm_100a -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 512

template <int V>
__global__ void k(unsigned* out, unsigned seed, long long* cyc) {
  unsigned r[CHAINS], t[CHAINS];
  for (int i = 0; i < CHAINS; ++i) { r[i] = seed + threadIdx.x * 7 + i; t[i] = seed * 3 + i; }
  unsigned one = 0x3c003c00u, c = seed | 0x3c003c01u;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (V == 0) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
      if (V == 1) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(one), "r"(c));
      if (V == 2) asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
      if (V == 3) asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(r[i]) : "r"(c));
      if (V == 4) {  // add + fma(x,1,y) alternating chains
        if (i & 1) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
        else asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(one), "r"(c));
      }
      if (V == 5) {  // mul + prmt
        if (i & 1) asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
        else asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(r[i]) : "r"(c));
      }
      if (V == 6) {  // add + prmt
        if (i & 1) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
        else asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(r[i]) : "r"(c));
      }
      if (V == 7) {  // mul + add
        if (i & 1) asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
        else asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c));
      }
      if (V == 8) {  // refine mix: per 6 -> 2 prmt, 2 mul, 1 add, 1 fma
        asm volatile("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(t[i]) : "r"(r[i]), "r"(c));
        asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(t[i]) : "r"(c));
        if (i & 1) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(t[i]));
        else asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(one), "r"(t[i]));
      }
      if (V == 9) {  // same, adds only via add
        asm volatile("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(t[i]) : "r"(r[i]), "r"(c));
        asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(t[i]) : "r"(c));
        asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(t[i]));
      }
      if (V == 10) {  // fp32 ffma for reference
        float f = __uint_as_float(r[i]);
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(1.0001f), "f"(0.5f));
        r[i] = __float_as_uint(f);
      }
    }
  }
  long long t1 = clock64();
  unsigned acc = 0;
  for (int i = 0; i < CHAINS; ++i) acc ^= r[i] ^ t[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int V>
void run(const char* name, int per_iter) {
  unsigned* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int warps_per_smsp = 8;
  k<V><<<148, 32 * 4 * warps_per_smsp>>>(out, 1, cyc);
  k<V><<<148, 32 * 4 * warps_per_smsp>>>(out, 1, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double instr = (double)ITERS * CHAINS * per_iter * warps_per_smsp;  // warp-instr per SMSP
  printf("%-34s %7.3f cycles / warp-instr / SMSP\n", name, h / instr);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("HADD2", 1);
  run<1>("HFMA2 (x*1+y)", 1);
  run<2>("HMUL2", 1);
  run<3>("PRMT", 1);
  run<4>("HADD2 + HFMA2 alternating", 1);
  run<5>("HMUL2 + PRMT alternating", 1);
  run<6>("HADD2 + PRMT alternating", 1);
  run<7>("HMUL2 + HADD2 alternating", 1);
  run<8>("mix prmt,mul,(add|fma)", 3);
  run<9>("mix prmt,mul,add", 3);
  run<10>("FFMA", 1);
  return 0;
}
