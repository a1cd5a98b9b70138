// Issue rates of fp32 scalar / packed FMA on sm_100a (cycles per warp-instruction per SM
// sub-partition, 8 warps per sub-partition, 8 independent chains per thread).
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 512

template <int V>
__global__ void k(float* out, float seed, long long* cyc) {
  float r[CHAINS], s[CHAINS];
  float2 p[CHAINS];
  for (int i = 0; i < CHAINS; ++i) {
    r[i] = seed + threadIdx.x * 0.001f + i;
    s[i] = seed * 0.5f + i;
    p[i] = make_float2(r[i], s[i]);
  }
  float b = seed * 1.0001f, c = seed * 0.37f;
  float2 b2 = make_float2(b, c), c2 = make_float2(c, b);
  unsigned x = threadIdx.x;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (V == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(r[i]) : "f"(b), "f"(c));
      if (V == 1) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(r[i]) : "f"(s[i]), "f"(b));
      if (V == 2) {
        unsigned long long pp, bb, cc;
        pp = *reinterpret_cast<unsigned long long*>(&p[i]);
        bb = *reinterpret_cast<unsigned long long*>(&b2);
        cc = *reinterpret_cast<unsigned long long*>(&c2);
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pp) : "l"(bb), "l"(cc));
        *reinterpret_cast<unsigned long long*>(&p[i]) = pp;
      }
      if (V == 3) {  // FFMA + FFMA2 alternating
        if (i & 1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(r[i]) : "f"(b), "f"(c));
        else {
          unsigned long long pp = *reinterpret_cast<unsigned long long*>(&p[i]);
          unsigned long long bb = *reinterpret_cast<unsigned long long*>(&b2);
          unsigned long long cc = *reinterpret_cast<unsigned long long*>(&c2);
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pp) : "l"(bb), "l"(cc));
          *reinterpret_cast<unsigned long long*>(&p[i]) = pp;
        }
      }
      if (V == 4) {  // FFMA + integer add (alu pipe)
        if (i & 1) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(r[i]) : "f"(b), "f"(c));
        else asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(i + 1));
      }
      if (V == 5) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(r[i]) : "f"(b));
      if (V == 6) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(r[i]) : "f"(b));
      if (V == 7) {  // FFMA2 + integer add
        if (i & 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(i + 1));
        else {
          unsigned long long pp = *reinterpret_cast<unsigned long long*>(&p[i]);
          unsigned long long bb = *reinterpret_cast<unsigned long long*>(&b2);
          unsigned long long cc = *reinterpret_cast<unsigned long long*>(&c2);
          asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(pp) : "l"(bb), "l"(cc));
          *reinterpret_cast<unsigned long long*>(&p[i]) = pp;
        }
      }
      if (V == 8) asm volatile("mov.b32 %0, %1;" : "=f"(s[i]) : "f"(r[(i + 1) % CHAINS]));
    }
  }
  long long t1 = clock64();
  float acc = __uint_as_float(x);
  for (int i = 0; i < CHAINS; ++i) acc += r[i] + s[i] + p[i].x + p[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int V>
void run(const char* name) {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int wps = 8;
  k<V><<<148, 32 * 4 * wps>>>(out, 1.0f, cyc);
  k<V><<<148, 32 * 4 * wps>>>(out, 1.0f, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-34s %7.3f cycles / warp-instr / SMSP\n", name, h / ((double)ITERS * CHAINS * wps));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("FFMA r=r*b+c");
  run<1>("FFMA r=s*b+r");
  run<2>("FFMA2");
  run<3>("FFMA + FFMA2 alternating");
  run<4>("FFMA + IADD alternating");
  run<5>("FMUL");
  run<6>("FADD");
  run<7>("FFMA2 + IADD alternating");
  run<8>("MOV");
  return 0;
}
