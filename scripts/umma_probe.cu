// Standalone probe (not part of the library): validates the no-swizzle K-major UMMA descriptor
// encoding used by the conv kernels against a CPU GEMM, including row-shifted A views, and
// measures tcgen05.mma issue cost for M=128, K=16 at several N (SMEM-operand bound or not?).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/umma_probe.cu -o scripts/umma_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include <cuda_fp16.h>
#include "../honk_b200/csrc/umma.cuh"

using namespace honk::umma;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int PROWS = 256;   // rows of the A planar buffer (128 + room for shifts)

// A planar: [KP planes][PROWS rows][8 halves]; B planar: [KP planes][N rows][8 halves]
__global__ void __launch_bounds__(128, 1) gemm_probe(const __half* __restrict__ gA, const __half* __restrict__ gB,
                                                     float* __restrict__ gD, int N, int KC, int shift, int swap) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int KP = KC * 2;
  const uint32_t planeA = PROWS * 16, planeB = N * 16;
  unsigned char* sA = smem;
  unsigned char* sB = smem + KP * planeA;
  for (int i = threadIdx.x; i < KP * PROWS * 8; i += blockDim.x) reinterpret_cast<__half*>(sA)[i] = gA[i];
  for (int i = threadIdx.x; i < KP * N * 8; i += blockDim.x) reinterpret_cast<__half*>(sB)[i] = gB[i];
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 256);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(128, N, 0);
    for (int kc = 0; kc < KC; ++kc) {
      const uint32_t a0 = smem_u32(sA) + kc * 2 * planeA + shift * 16;
      const uint32_t b0 = smem_u32(sB) + kc * 2 * planeB;
      const uint64_t ad = swap ? make_desc(a0, 128, planeA) : make_desc(a0, planeA, 128);
      const uint64_t bd = swap ? make_desc(b0, 128, planeB) : make_desc(b0, planeB, 128);
      mma_f16_ss(tmem, ad, bd, idesc, kc > 0);
    }
    commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  fence_after();
  for (int c0 = 0; c0 < N; c0 += 8) {
    uint32_t v[8];
    tmem_ld8(tmem + ((warp * 32u) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 8; ++j) gD[threadIdx.x * N + c0 + j] = __uint_as_float(v[j]);
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// throughput: `iters` rounds of 9 "taps" x KC MMAs with rotating A shifts, like the conv main loop
__global__ void __launch_bounds__(128, 1) rate_probe(long long* out, int N, int KC, int iters, int same_a) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int KP = KC * 2;
  const uint32_t planeA = 512 * 16, planeB = 9 * N * 16;
  unsigned char* sA = smem;
  unsigned char* sB = smem + KP * planeA;
  for (int i = threadIdx.x; i < (KP * (int)planeA + KP * (int)planeB) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;   // 1.0h
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(128, N, 0);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t acc = tmem + (it & 1) * 256;
      for (int tap = 0; tap < 9; ++tap) {
        const int shift = same_a ? 0 : ((tap / 3) * 41 + (tap % 3)) + (it & 3) * 64;
        for (int kc = 0; kc < KC; ++kc) {
          const uint64_t ad = make_desc(smem_u32(sA) + kc * 2 * planeA + shift * 16, planeA, 128);
          const uint64_t bd = make_desc(smem_u32(sB) + kc * 2 * planeB + tap * N * 16, planeB, 128);
          mma_f16_ss(acc, ad, bd, idesc, (tap | kc) > 0);
        }
      }
    }
    commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    const long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}

// variant 2: warp-uniform issue (elect.sync), descriptors precomputed; mode 0 = no-swizzle planar,
// mode 1 = 128B-swizzle K-major rows (timing only), M = 128 or 64
__global__ void __launch_bounds__(128, 1) rate_probe2(long long* out, int N, int M, int mode, int iters) {
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
    const uint32_t idesc = make_idesc(M, N, 0);
    uint64_t ad[12], bd[12];
    const uint32_t sA = smem_u32(smem), sB = smem_u32(smem) + 100 * 1024;
    for (int j = 0; j < 12; ++j) {
      if (mode == 0) {
        ad[j] = make_desc(sA + (j % 3) * 2 * 8192 + (j / 3) * 41 * 16, 8192, 128);
        bd[j] = make_desc(sB + j * 2 * N * 16, N * 16, 128);
      } else {
        ad[j] = make_desc(sA + (j / 4) * 16384 + (j % 4) * 32, 16, 1024) | (2ull << 61);
        bd[j] = make_desc(sB + (j / 4) * 32768 + (j % 4) * 32, 16, 1024) | (2ull << 61);
      }
    }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t acc = tmem + (it & 1) * 256;
      if (elect_one()) {
#pragma unroll
        for (int j = 0; j < 12; ++j) mma_f16_ss(acc, ad[j], bd[j], idesc, j > 0);
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
  // ---------------- correctness ----------------
  const int N = 48, KC = 3, KP = KC * 2, K = KC * 16;
  std::vector<__half> hA(KP * PROWS * 8), hB(KP * N * 8);
  std::vector<float> fA(PROWS * K), fB(N * K);
  srand(1);
  for (int r = 0; r < PROWS; ++r) for (int k = 0; k < K; ++k) {
    float v = (rand() % 17 - 8) / 8.0f; fA[r * K + k] = v; hA[((k / 8) * PROWS + r) * 8 + k % 8] = __float2half(v);
  }
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) {
    float v = (rand() % 13 - 6) / 4.0f; fB[n * K + k] = v; hB[((k / 8) * N + n) * 8 + k % 8] = __float2half(v);
  }
  __half *dA, *dB; float* dD;
  CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2)); CK(cudaMalloc(&dD, 128 * N * 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
  const int smem_g = KP * PROWS * 16 + KP * N * 16;
  CK(cudaFuncSetAttribute(gemm_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_g));
  for (int swap = 0; swap < 1; ++swap) for (int shift : {0, 5, 41}) {
    CK(cudaMemset(dD, 0, 128 * N * 4));
    gemm_probe<<<1, 128, smem_g>>>(dA, dB, dD, N, KC, shift, swap);
    CK(cudaDeviceSynchronize());
    std::vector<float> hD(128 * N);
    CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0;
    for (int m = 0; m < 128; ++m) for (int n = 0; n < N; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fA[(m + shift) * K + k] * fB[n * K + k];
      maxerr = fmax(maxerr, fabs(ref - hD[m * N + n]));
    }
    printf("gemm_probe swap=%d shift=%d max_err=%g %s\n", swap, shift, maxerr, maxerr < 1e-3 ? "OK" : "MISMATCH");
  }
  // ---------------- throughput ----------------
  long long* dT; CK(cudaMalloc(&dT, 148 * 8));
  for (int grid : {1, 148}) for (int same_a : {0, 1}) for (int n : {48, 96, 144, 192, 256}) {
    const int kc = 3, iters = 200;
    const int smem_r = kc * 2 * 512 * 16 + kc * 2 * 9 * n * 16;
    if (smem_r > 227 * 1024) continue;
    CK(cudaFuncSetAttribute(rate_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_r));
    rate_probe<<<grid, 128, smem_r>>>(dT, n, kc, iters, same_a);
    CK(cudaDeviceSynchronize());
    std::vector<long long> hT(grid);
    CK(cudaMemcpy(hT.data(), dT, grid * 8, cudaMemcpyDeviceToHost));
    long long mx = 0; for (auto t : hT) mx = t > mx ? t : mx;
    const double per = (double)mx / (iters * 9 * kc);
    printf("rate_probe grid=%3d same_a=%d N=%3d: %.1f clk/MMA  (floor N/2=%.0f)  -> %.0f flop/clk/SM\n", grid, same_a, n,
           per, n / 2.0, 2.0 * 128 * n * 16 / per);
  }
  CK(cudaFuncSetAttribute(rate_probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (int mode : {0, 1}) for (int m : {128, 64}) for (int n : {48, 96, 144, 256}) {
    const int iters = 500;
    rate_probe2<<<1, 128, 200 * 1024>>>(dT, n, m, mode, iters);
    CK(cudaDeviceSynchronize());
    long long t; CK(cudaMemcpy(&t, dT, 8, cudaMemcpyDeviceToHost));
    const double per = (double)t / (iters * 12);
    printf("rate_probe2 mode=%d M=%3d N=%3d: %.1f clk/MMA -> %.0f flop/clk/SM\n", mode, m, n, per, 2.0 * m * n * 16 / per);
  }
  return 0;
}
