// Micro-benchmark (development aid, not part of the product): FP64 issue rates on one SM-full grid.
//   mode 0: DFMA chains (26 independent accumulators per lane)
//   mode 1: mma.sync m8n8k4 f64 (16 independent accumulator fragments)
//   mode 2: mma.sync m16n8k4 f64      mode 3: m16n8k8      mode 4: m16n8k16
// Prints cycles per instruction per warp scheduler and the implied FMA/clk/SM.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(128, 1) probe(double* out, int iters, long long* cyc) {
  const int lane = threadIdx.x & 31;
  double a0 = 1.0 + lane * 1e-9, b0 = 1.0 - lane * 1e-9;
  double acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 26; ++i) acc[i] = fma(a0, b0, acc[i]);
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(acc[2 * i]), "+d"(acc[2 * i + 1]) : "d"(a0), "d"(b0));
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(acc[4 * i]), "+d"(acc[4 * i + 1]), "+d"(acc[4 * i + 2]), "+d"(acc[4 * i + 3])
                     : "d"(a0), "d"(b0), "d"(b0));
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(acc[4 * i]), "+d"(acc[4 * i + 1]), "+d"(acc[4 * i + 2]), "+d"(acc[4 * i + 3])
                     : "d"(a0), "d"(b0), "d"(a0), "d"(b0), "d"(b0), "d"(a0));
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                     : "+d"(acc[4 * i]), "+d"(acc[4 * i + 1]), "+d"(acc[4 * i + 2]), "+d"(acc[4 * i + 3])
                     : "d"(a0), "d"(b0), "d"(a0), "d"(b0), "d"(a0), "d"(b0), "d"(a0), "d"(b0), "d"(b0), "d"(a0), "d"(b0), "d"(a0));
    }
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// shared-memory broadcast cost: every lane reads the same address; WIDTH 8 or 16 bytes; LANES_DISTINCT variant for reference
template <int WIDTH, bool BCAST>
__global__ void __launch_bounds__(128, 1) lds_probe(double* out, int iters, long long* cyc) {
  __shared__ __align__(16) double buf[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) buf[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  double s0 = 0, s1 = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = (BCAST ? 0 : lane * (WIDTH / 8)) + ((it + i) & 7) * 64 * (WIDTH / 8);
      if (WIDTH == 16) {
        double vx, vy;
        asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(vx), "=d"(vy) : "r"(static_cast<unsigned>(__cvta_generic_to_shared(&buf[idx & 4094]))));
        s0 += vx; s1 += vy;
      } else {
        double vx;
        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(vx) : "r"(static_cast<unsigned>(__cvta_generic_to_shared(&buf[idx & 4095]))));
        s0 += vx;
      }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 148 * 128 * 8); cudaMallocManaged(&cyc, 8);
  const int iters = 20000;
  const char* names[] = {"DFMA x26", "DMMA m8n8k4 x16", "DMMA m16n8k4 x8", "DMMA m16n8k8 x8", "DMMA m16n8k16 x8"};
  const double fma_per_inst[] = {32, 256, 512, 1024, 2048};
  const int inst_per_iter[] = {26, 16, 8, 8, 8};
  for (int m = 0; m < 5; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      if (m == 0) probe<0><<<148, 128>>>(out, iters, cyc);
      if (m == 1) probe<1><<<148, 128>>>(out, iters, cyc);
      if (m == 2) probe<2><<<148, 128>>>(out, iters, cyc);
      if (m == 3) probe<3><<<148, 128>>>(out, iters, cyc);
      if (m == 4) probe<4><<<148, 128>>>(out, iters, cyc);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: %s\n", names[m], cudaGetErrorString(e)); return 1; }
    }
    const double cpi = double(*cyc) / (double(iters) * inst_per_iter[m]);
    printf("%-20s %.2f clk/inst/scheduler  -> %.1f FMA/clk/SM\n", names[m], cpi, 4 * fma_per_inst[m] / cpi);
  }
  for (int m = 0; m < 4; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      if (m == 0) lds_probe<16, true><<<148, 128>>>(out, iters, cyc);
      if (m == 1) lds_probe<8, true><<<148, 128>>>(out, iters, cyc);
      if (m == 2) lds_probe<16, false><<<148, 128>>>(out, iters, cyc);
      if (m == 3) lds_probe<8, false><<<148, 128>>>(out, iters, cyc);
      cudaDeviceSynchronize();
    }
    const char* n2[] = {"LDS.128 broadcast", "LDS.64 broadcast", "LDS.128 distinct", "LDS.64 distinct"};
    printf("%-20s %.2f clk/inst/SM (4 warps)\n", n2[m], double(*cyc) / (double(iters) * 16 * 4));
  }
  return 0;
}
