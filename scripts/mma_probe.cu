// Probe 3: which property of the ring-accumulate issue pattern slows tcgen05.mma down?
// Variants: A K-half stride (LBO) 8192 vs 1872 bytes, D column base pattern, N.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include "../honk_b200/csrc/umma.cuh"
using namespace honk::umma;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
__global__ void __launch_bounds__(128, 1) probe(long long* out, int N, int lbo_rows, int dmode, int d, int iters) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (warp == 1) {
    const uint32_t idesc = make_idesc(128, N, 0);
    const uint32_t sA = smem_u32(smem), sB = smem_u32(smem) + 100 * 1024;
    const uint64_t a_hi = make_desc(0, lbo_rows * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a_lbo = (static_cast<uint32_t>(lbo_rows) & 0x3FFF) << 16;
    const uint64_t b_hi = make_desc(0, N * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t b_lbo = (static_cast<uint32_t>(N) & 0x3FFF) << 16;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      uint32_t dcol;
      if (dmode == 0) dcol = (it & 1) * 256;
      else if (dmode == 1) dcol = (it % 7) * 48;
      else dcol = (it % 3) * 144;
      const uint32_t a_lo0 = a_lbo | ((sA + (it & 7) * 12032) >> 4);
      const uint32_t b_lo0 = b_lbo | (sB >> 4);
      if (elect_one()) {
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const int kh = i / 3, kc = i % 3;
          const uint64_t ad = a_hi | static_cast<uint64_t>(a_lo0 + static_cast<uint32_t>(2 * kc * lbo_rows + kh * d));
          const uint64_t bd = b_hi | static_cast<uint64_t>(b_lo0 + static_cast<uint32_t>(i * 2 * N));
          mma_f16_ss(tmem + dcol, ad, bd, idesc, i > 0);
        }
      }
      __syncwarp();
    }
    if (elect_one()) commit(smem_u32(&bar));
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) out[blockIdx.x] = t1 - t0;
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}
int main() {
  long long* dT; CK(cudaMalloc(&dT, 148 * 8));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  const int iters = 400;
  for (int N : {144, 48}) for (int lbo : {512, 128, 117, 120, 136}) for (int dmode : {0, 1, 2}) for (int d : {0, 1, 16}) {
    if (N == 48 && dmode == 2) continue;
    probe<<<1, 128, 200 * 1024>>>(dT, N, lbo, dmode, d, iters);
    CK(cudaDeviceSynchronize());
    long long t; CK(cudaMemcpy(&t, dT, 8, cudaMemcpyDeviceToHost));
    printf("mma_probe N=%3d lbo_rows=%3d dmode=%d d=%2d: %.1f clk/MMA, %.0f clk/task\n", N, lbo, dmode, d, (double)t / (iters * 9), (double)t / iters);
  }
  return 0;
}
