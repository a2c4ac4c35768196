// Development aid: cost of mixing DFMA / DMUL with DMMA.8x8x4 in one warp's instruction stream (1 warp per scheduler).
//   mode 0: 16 DMMA                       mode 1: 16 DMMA, then 16 dependent-pair DMUL/DFMA (grouped)
//   mode 2: DMMA and DFMA interleaved 1:1  mode 3: 16 DMMA + 16 independent DFMA grouped
//   mode 4: as 1 but a second warp per scheduler runs an independent DFMA stream (foreign traffic)
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
template <int MODE>
__global__ void __launch_bounds__(256, 1) probe(double* out, int iters, long long* cyc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double a0 = 1.0 + lane * 1e-9, b0 = 1.0 - lane * 1e-9;
  double acc[32], c[4], s[4];
  const double rc = 0.999, rs = 0.04;
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.0;
#pragma unroll
  for (int i = 0; i < 4; ++i) { c[i] = 1.0 - i * 1e-3; s[i] = i * 1e-3; }
  __syncthreads();
  const long long t0 = clock64();
  if (warp < 4) {
    for (int it = 0; it < iters; ++it) {
      if (MODE == 2) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          dmma(acc[2 * i], acc[2 * i + 1], a0, c[i & 3]);
          asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(s[i & 3]) : "d"(a0), "d"(b0));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) dmma(acc[2 * i], acc[2 * i + 1], a0, (MODE == 1 || MODE == 4) ? ((i & 1) ? s[(i >> 1) & 3] : c[(i >> 1) & 3]) : b0);
        if (MODE == 1 || MODE == 4) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const double nc = fma(c[i], rc, -s[i] * rs);
            s[i] = fma(s[i], rc, c[i] * rs);
            c[i] = nc;
          }
        }
        if (MODE == 3) {
#pragma unroll
          for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(acc[(2 * i) & 31]) : "d"(a0), "d"(b0));
        }
      }
    }
  } else if (MODE == 4) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 4; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(acc[i]) : "d"(a0), "d"(b0));
    }
  }
  const long long t1 = clock64();
  double r = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) r += acc[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) r += c[i] + s[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 148 * 256 * 8); cudaMallocManaged(&cyc, 8);
  const int iters = 20000;
  const char* names[] = {"16 DMMA", "16 DMMA + 4 rotations (16 FP64) grouped", "16 x (DMMA, DFMA) interleaved", "16 DMMA + 16 indep DFMA grouped", "mode 1 + foreign warp 4 DFMA/iter"};
  for (int m = 0; m < 5; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      if (m == 0) probe<0><<<148, 256>>>(out, iters, cyc);
      if (m == 1) probe<1><<<148, 256>>>(out, iters, cyc);
      if (m == 2) probe<2><<<148, 256>>>(out, iters, cyc);
      if (m == 3) probe<3><<<148, 256>>>(out, iters, cyc);
      if (m == 4) probe<4><<<148, 256>>>(out, iters, cyc);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
    }
    printf("%-45s %.1f clk per iteration (warp 0)\n", names[m], double(*cyc) / iters);
  }
  return 0;
}
