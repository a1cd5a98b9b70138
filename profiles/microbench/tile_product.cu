// Microbenchmark: P = A B^T for two 32x32 fp64 tiles in shared memory by one 256-thread CTA.
//   v0: the kernel's tile_product (4x4 register blocks, inner dimension split over four
//       64-thread groups, partial tiles reduced through shared memory)
//   v1: mma.sync.m8n8k4.f64 (DMMA): warp w computes rows 8(w&3).., columns 16(w>>2).. over the whole
//       inner dimension (16 DMMA per warp), result written to a shared tile and re-read in the
//       kernel's register layout (elements 4t..4t+3)
// Prints clocks per product and checks v1 against v0.
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tile_product tile_product.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
constexpr int TB = 32, kTile = 1024, kCT = 256;
__host__ __device__ inline int swz(int r, int c) {
  return (c << 5) + ((((r >> 1) ^ ((c >> 1) & 7)) << 1) | (r & 1));
}
__device__ __forceinline__ void tile_mma(const double* __restrict__ A, const double* __restrict__ B,
                                         int ty, int tx, int k0, int k1, double (&acc)[4][4]) {
#pragma unroll 8
  for (int kk = k0; kk < k1; ++kk) {
    const int s = (kk >> 1) & 7;
    const double2 a0 = *reinterpret_cast<const double2*>(A + (kk << 5) + ((ty ^ s) << 1));
    const double2 a1 = *reinterpret_cast<const double2*>(A + (kk << 5) + (((ty + 8) ^ s) << 1));
    const double2 b0 = *reinterpret_cast<const double2*>(B + (kk << 5) + ((tx ^ s) << 1));
    const double2 b1 = *reinterpret_cast<const double2*>(B + (kk << 5) + (((tx + 8) ^ s) << 1));
    const double av[4] = {a0.x, a0.y, a1.x, a1.y};
    const double bv[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
  }
}
__device__ __forceinline__ int out_pair_off(int ty, int tx, int a, int b) {
  const int c = 2 * tx + (b & 1) + 16 * (b >> 1);
  return (c << 5) + (((ty + 8 * (a >> 1)) ^ tx) << 1);
}
__device__ __forceinline__ void product_v0(const double* A, const double* B, double* red, int tid,
                                           double (&out)[4]) {
  const int g = tid >> 6, tg = tid & 63, ty = tg >> 3, tx = tg & 7;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  tile_mma(A, B, ty, tx, 8 * g, 8 * g + 8, acc);
  double* mine = red + g * kTile;
#pragma unroll
  for (int b = 0; b < 4; ++b)
#pragma unroll
    for (int a = 0; a < 4; a += 2)
      *reinterpret_cast<double2*>(mine + out_pair_off(ty, tx, a, b)) = make_double2(acc[a][b], acc[a + 1][b]);
  __syncthreads();
  const double2* r2 = reinterpret_cast<const double2*>(red);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int q = 2 * tid + h;
    const double2 p0 = r2[q], p1 = r2[512 + q], p2 = r2[1024 + q], p3 = r2[1536 + q];
    out[2 * h] = (p0.x + p1.x) + (p2.x + p3.x);
    out[2 * h + 1] = (p0.y + p1.y) + (p2.y + p3.y);
  }
  __syncthreads();
}
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void product_v1(const double* A, const double* B, double* red, int tid,
                                           double (&out)[4]) {
  const int warp = tid >> 5, lane = tid & 31;
  const int r0 = 8 * (warp & 3), c0 = 16 * (warp >> 2);
  const int g = lane >> 2, q = lane & 3;
  double c00 = 0, c01 = 0, c10 = 0, c11 = 0;
#pragma unroll
  for (int k0 = 0; k0 < TB; k0 += 4) {
    const double a = A[swz(r0 + g, k0 + q)];
    const double b0 = B[swz(c0 + g, k0 + q)];
    const double b1 = B[swz(c0 + 8 + g, k0 + q)];
    dmma(c00, c01, a, b0);
    dmma(c10, c11, a, b1);
  }
  red[swz(r0 + g, c0 + 2 * q)] = c00;
  red[swz(r0 + g, c0 + 2 * q + 1)] = c01;
  red[swz(r0 + g, c0 + 8 + 2 * q)] = c10;
  red[swz(r0 + g, c0 + 8 + 2 * q + 1)] = c11;
  __syncthreads();
  const double2* r2 = reinterpret_cast<const double2*>(red);
  const double2 p0 = r2[2 * tid], p1 = r2[2 * tid + 1];
  out[0] = p0.x; out[1] = p0.y; out[2] = p1.x; out[3] = p1.y;
  __syncthreads();
}
template <int V>
__global__ void __launch_bounds__(kCT, 1) bench(const double* gA, const double* gB, double* gP,
                                                long long* clocks, int reps) {
  extern __shared__ __align__(16) double sm[];
  double* A = sm; double* B = sm + kTile; double* red = sm + 2 * kTile;
  const int tid = threadIdx.x;
  for (int e = tid; e < kTile; e += kCT) { A[e] = gA[e]; B[e] = gB[e]; }
  __syncthreads();
  double out[4] = {0, 0, 0, 0}, sum[4] = {0, 0, 0, 0};
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    if (V == 0) product_v0(A, B, red, tid, out); else product_v1(A, B, red, tid, out);
#pragma unroll
    for (int q = 0; q < 4; ++q) sum[q] += out[q];
  }
  const long long t1 = clock64();
  if (tid == 0) clocks[V] = (t1 - t0) / reps;
#pragma unroll
  for (int q = 0; q < 4; ++q) gP[V * kTile + 4 * tid + q] = sum[q] / reps;
}
int main() {
  double *hA = new double[kTile], *hB = new double[kTile];
  srand(1);
  for (int r = 0; r < TB; ++r) for (int c = 0; c < TB; ++c) {
    hA[swz(r, c)] = (rand() % 2001 - 1000) / 1000.0; hB[swz(r, c)] = (rand() % 2001 - 1000) / 1000.0; }
  double *dA, *dB, *dP; long long* dC;
  cudaMalloc(&dA, kTile * 8); cudaMalloc(&dB, kTile * 8); cudaMalloc(&dP, 2 * kTile * 8); cudaMalloc(&dC, 16);
  cudaMemcpy(dA, hA, kTile * 8, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, kTile * 8, cudaMemcpyHostToDevice);
  const size_t smem = 6 * kTile * 8;
  cudaFuncSetAttribute(bench<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(bench<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int it = 0; it < 2; ++it) {
    bench<0><<<1, kCT, smem>>>(dA, dB, dP, dC, 200);
    bench<1><<<1, kCT, smem>>>(dA, dB, dP, dC, 200);
  }
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("error %s\n", cudaGetErrorString(err)); return 1; }
  long long hC[2]; double* hP = new double[2 * kTile];
  cudaMemcpy(hC, dC, 16, cudaMemcpyDeviceToHost); cudaMemcpy(hP, dP, 2 * kTile * 8, cudaMemcpyDeviceToHost);
  double maxd = 0, maxref = 0;
  for (int r = 0; r < TB; ++r) for (int c = 0; c < TB; ++c) {
    double ref = 0; for (int k = 0; k < TB; ++k) ref += hA[swz(r, k)] * hB[swz(c, k)];
    maxref = fmax(maxref, fabs(hP[swz(r, c)] - ref));
    maxd = fmax(maxd, fabs(hP[swz(r, c)] - hP[kTile + swz(r, c)])); }
  printf("v0 (register blocks + reduction): %lld clocks per product (%.3f us at 1965 MHz), max |v0 - host| %.2e\n",
         hC[0], hC[0] / 1965.0, maxref);
  printf("v1 (DMMA m8n8k4):                 %lld clocks per product (%.3f us), max |v1 - v0| %.2e\n",
         hC[1], hC[1] / 1965.0, maxd);
  return 0;
}
