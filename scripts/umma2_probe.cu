// Standalone probe (not part of the library): tcgen05.mma.cta_group::2 with the conv kernels' no-swizzle K-major
// operands.  (1) correctness of the pair semantics the planned 2-CTA conv kernel relies on: M = 256 = 128 rows of A
// from each CTA's shared memory at the SAME descriptor offset, B split along N (first N/2 rows from CTA 0, second N/2
// from CTA 1, again at the same offset), D rows r*128.. in CTA r's TMEM; (2) MMA rate for M128xN144xK16 (single CTA)
// vs M256xN144xK16 (pair) with and without a concurrent cp.async.bulk stream of one 12 KB column per 9 MMAs into the
// same shared memory - the question being whether halving the B operand reads per SM takes the MMA off the
// shared-memory port limit.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/umma2_probe.cu -o scripts/umma2_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include <cuda_fp16.h>
#include "../honk_b200/csrc/umma.cuh"

using namespace honk::umma;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mma2_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void commit2_mc(uint32_t mbar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(mbar),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(mbar)
               : "memory");
}

// ---- (1) correctness ---------------------------------------------------------------------------
// gA: [2 CTAs][2 K halves][128 rows][8]; gB: [2 CTAs][2 K halves][N/2 rows][8]; gD: [256][N]
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
    gemm2_probe(const __half* __restrict__ gA, const __half* __restrict__ gB, float* __restrict__ gD, int N) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const uint32_t rank = cluster_ctarank();
  const int Nh = N / 2;
  unsigned char* sA = smem;                 // 2 * 128 * 16 = 4096 B
  unsigned char* sB = smem + 4096;          // 2 * Nh * 16
  for (int i = threadIdx.x; i < 2 * 128 * 8; i += blockDim.x) reinterpret_cast<__half*>(sA)[i] = gA[rank * 2 * 128 * 8 + i];
  for (int i = threadIdx.x; i < 2 * Nh * 8; i += blockDim.x) reinterpret_cast<__half*>(sB)[i] = gB[rank * 2 * Nh * 8 + i];
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc2(smem_u32(&tmem_base_s), 256);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  cluster_sync();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (rank == 0 && threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(256, N, 0);
    const uint64_t ad = make_desc(smem_u32(sA), 128 * 16, 128);
    const uint64_t bd = make_desc(smem_u32(sB), Nh * 16, 128);
    mma2_f16_ss(tmem, ad, bd, idesc, 0);
    commit2_mc(smem_u32(&bar), 3);
  }
  mbar_wait(smem_u32(&bar), 0);
  fence_after();
  for (int c0 = 0; c0 < N; c0 += 8) {
    uint32_t v[8];
    tmem_ld8(tmem + ((warp * 32u) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 8; ++j) gD[(rank * 128 + threadIdx.x) * N + c0 + j] = __uint_as_float(v[j]);
  }
  fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 0) tmem_dealloc2(tmem, 256);
}

// ---- (2) rate ------------------------------------------------------------------------------------
// `iters` tasks of 9 MMAs (3 row shifts x 3 K chunks, 9 distinct B tiles) like one input column of the conv kernel.
// PAIR: cluster of 2, leader issues cta_group::2 MMAs (B tiles hold N/2 rows per CTA).  copies: thread 64 of every CTA
// keeps a ring of four 12 KB bulk copies from an L2-resident buffer in flight while the MMAs run.
template <bool PAIR>
__global__ void __launch_bounds__(128, 1) rate2_probe(long long* out, const unsigned char* __restrict__ gsrc, int N, int iters, int copies,
                                                      int shift_rows = 16, int plane_rows = 160) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar, cbar[4];
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int stop;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  const int NB = PAIR ? N / 2 : N;
  const int PLANE = plane_rows * 16;                    // A: 6 planes (3 K chunks x 2 halves) of plane_rows rows (room: 6 x 2560 B)
  unsigned char* sA = smem;                             // 6 * 2560 = 15360
  unsigned char* sB = smem + 6 * 2560;                  // 9 tiles * 2 * NB * 16
  unsigned char* sC = sB + 9 * 2 * NB * 16;             // copy ring: 4 x 12288
  const int words = (6 * 2560 + 9 * 2 * NB * 16) / 4;
  for (int i = threadIdx.x; i < words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { if (PAIR) tmem_alloc2(smem_u32(&tmem_base_s), 256); else tmem_alloc(smem_u32(&tmem_base_s), 256); }
  if (threadIdx.x == 32) {
    mbar_init(smem_u32(&bar), 1);
    for (int k = 0; k < 4; ++k) mbar_init(smem_u32(&cbar[k]), 1);
    fence_mbar_init();
    stop = 0;
  }
  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  long long n_copies = 0;
  if (threadIdx.x == 64 && copies) {
    uint32_t it = 0;
    for (int k = 0; k < 4; ++k, ++it) {
      mbar_expect_tx(smem_u32(&cbar[k]), 12288);
      bulk_load(smem_u32(sC + k * 12288), gsrc + (it % 64) * 12288, 12288, smem_u32(&cbar[k]));
    }
    while (!stop) {
      const int k = it & 3;
      mbar_wait(smem_u32(&cbar[k]), ((it >> 2) & 1) ^ 1);
      ++n_copies;
      mbar_expect_tx(smem_u32(&cbar[k]), 12288);
      bulk_load(smem_u32(sC + k * 12288), gsrc + (it % 64) * 12288, 12288, smem_u32(&cbar[k]));
      ++it;
    }
    for (int k = 0; k < 4; ++k, ++it) mbar_wait(smem_u32(&cbar[it & 3]), ((it >> 2) & 1) ^ 1);
  }
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc = make_idesc(PAIR ? 256 : 128, N, 0);
    const long long t0 = clock64();
    for (int r = 0; r < iters; ++r) {
      for (int i = 0; i < 9; ++i) {
        const int kh = i / 3, kc = i - kh * 3;
        const uint64_t ad = make_desc(smem_u32(sA) + kc * 2 * PLANE + kh * shift_rows * 16, PLANE, 128);
        const uint64_t bd = make_desc(smem_u32(sB) + i * 2 * NB * 16, NB * 16, 128);
        if (PAIR) mma2_f16_ss(tmem, ad, bd, idesc, 1); else mma_f16_ss(tmem, ad, bd, idesc, 1);
      }
    }
    if (PAIR) commit2_mc(smem_u32(&bar), 3); else commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    out[0] = clock64() - t0;
  } else if (threadIdx.x == 0) {
    mbar_wait(smem_u32(&bar), 0);                         // follower: the pair's MMAs are done
  }
  if (threadIdx.x == 0) stop = 1;
  __syncthreads();
  if (threadIdx.x == 64) out[1 + rank] = n_copies;
  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync();
  if (warp == 0) { if (PAIR) tmem_dealloc2(tmem, 256); else tmem_dealloc(tmem, 256); }
}

// (3) kernel-like issue pattern: per task 9 MMAs into a rotating accumulator slot (48-column steps over 8 slots) followed by
// `ncommit` tcgen05.commit to rotating mbarriers nobody waits on.  Does the per-task bookkeeping cost tensor time?
__global__ void __launch_bounds__(128, 1) task_probe(long long* out, int N, int iters, int ncommit, int rotate) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar, cb[16];
  __shared__ uint32_t tmem_base_s;
  constexpr int PLANE = 160 * 16;
  unsigned char* sA = smem;
  unsigned char* sB = smem + 6 * PLANE;
  const int words = (6 * PLANE + 9 * 2 * 144 * 16) / 4;
  for (int i = threadIdx.x; i < words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (threadIdx.x == 32) {
    mbar_init(smem_u32(&bar), 1);
    for (int k = 0; k < 16; ++k) mbar_init(smem_u32(&cb[k]), 1);
    fence_mbar_init();
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(128, N, 0);
    const long long t0 = clock64();
    for (int r = 0; r < iters; ++r) {
      const uint32_t d = tmem + (rotate ? (r & 7) * 48 : 0);
      for (int i = 0; i < 9; ++i) {
        const int kh = i / 3, kc = i - kh * 3;
        const uint64_t ad = make_desc(smem_u32(sA) + kc * 2 * PLANE + kh * 16, PLANE, 128);
        const uint64_t bd = make_desc(smem_u32(sB) + i * 2 * 144 * 16, 144 * 16, 128);
        mma_f16_ss(d, ad, bd, idesc, 1);
      }
      for (int c = 0; c < ncommit; ++c) commit(smem_u32(&cb[(r * 2 + c) & 15]));
    }
    commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    out[0] = clock64() - t0;
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// (4) is the ~54 clk issue cost of a tcgen05.mma per issuing THREAD or per SM?  `nthr` threads (lane 0 of warps 0..nthr-1)
// each issue iters x 9 MMAs of width N into their own accumulator columns.
__global__ void __launch_bounds__(128, 1) dual_probe(long long* out, int N, int iters, int nthr) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar[4];
  __shared__ uint32_t tmem_base_s;
  constexpr int PLANE = 160 * 16;
  unsigned char* sA = smem;
  unsigned char* sB = smem + 6 * PLANE;
  const int words = (6 * PLANE + 9 * 2 * 144 * 16) / 4;
  for (int i = threadIdx.x; i < words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (threadIdx.x == 32) {
    for (int k = 0; k < 4; ++k) mbar_init(smem_u32(&bar[k]), 1);
    fence_mbar_init();
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const long long t0 = clock64();
  if ((threadIdx.x & 31) == 0 && warp < nthr) {
    const uint32_t idesc = make_idesc(128, N, 0);
    const uint32_t d = tmem + warp * 128;
    for (int r = 0; r < iters; ++r)
      for (int i = 0; i < 9; ++i) {
        const int kh = i / 3, kc = i - kh * 3;
        const uint64_t ad = make_desc(smem_u32(sA) + kc * 2 * PLANE + kh * 16, PLANE, 128);
        const uint64_t bd = make_desc(smem_u32(sB) + i * 2 * 144 * 16, 144 * 16, 128);
        mma_f16_ss(d, ad, bd, idesc, 1);
      }
    commit(smem_u32(&bar[warp]));
    mbar_wait(smem_u32(&bar[warp]), 0);
    out[warp] = clock64() - t0;
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  // ---- correctness ----
  for (int N : {144, 96, 48}) {
    const int Nh = N / 2;
    std::vector<__half> hA(2 * 2 * 128 * 8), hB(2 * 2 * Nh * 8);
    std::vector<float> fA(256 * 16), fB(N * 16);
    for (int r = 0; r < 2; ++r)
      for (int i = 0; i < 128; ++i)
        for (int k = 0; k < 16; ++k) {
          const float v = static_cast<float>(((r * 128 + i) * 7 + k * 3) % 11 - 5);
          fA[(r * 128 + i) * 16 + k] = v;
          hA[((r * 2 + k / 8) * 128 + i) * 8 + k % 8] = __float2half(v);
        }
    for (int n = 0; n < N; ++n)
      for (int k = 0; k < 16; ++k) {
        const float v = static_cast<float>((n * 5 + k * 2) % 7 - 3);
        fB[n * 16 + k] = v;
        const int r = n / Nh, nn = n % Nh;
        hB[((r * 2 + k / 8) * Nh + nn) * 8 + k % 8] = __float2half(v);
      }
    __half *dA, *dB;
    float* dD;
    CK(cudaMalloc(&dA, hA.size() * 2));
    CK(cudaMalloc(&dB, hB.size() * 2));
    CK(cudaMalloc(&dD, 256 * N * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xFF, 256 * N * 4));
    gemm2_probe<<<2, 128, 4096 + 2 * Nh * 16 + 1024>>>(dA, dB, dD, N);
    CK(cudaDeviceSynchronize());
    std::vector<float> hD(256 * N);
    CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
    double max_err = 0;
    for (int m = 0; m < 256; ++m)
      for (int n = 0; n < N; ++n) {
        float ref = 0;
        for (int k = 0; k < 16; ++k) ref += fA[m * 16 + k] * fB[n * 16 + k];
        max_err = std::fmax(max_err, std::fabs(ref - hD[m * N + n]));
      }
    printf("gemm2_probe M=256 N=%3d: max_err=%g %s   (A rows r*128.. from CTA r, B rows r*N/2.. from CTA r, same offsets)\n", N, max_err,
           max_err == 0 ? "OK" : "MISMATCH");
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
  }
  // ---- rate ----
  long long* dout;
  unsigned char* dsrc;
  CK(cudaMalloc(&dout, 64));
  CK(cudaMalloc(&dsrc, 64 * 12288));
  CK(cudaMemset(dsrc, 0, 64 * 12288));
  const int N = 144, iters = 2000;
  const int smem1 = 6 * 2560 + 9 * 2 * N * 16 + 4 * 12288 + 1024, smem2 = 6 * 2560 + 9 * 2 * (N / 2) * 16 + 4 * 12288 + 1024;
  CK(cudaFuncSetAttribute(rate2_probe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem1));
  CK(cudaFuncSetAttribute(rate2_probe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
  for (int copies = 0; copies < 2; ++copies) {
    long long h[3];
    CK(cudaMemset(dout, 0, 64));
    rate2_probe<false><<<1, 128, smem1>>>(dout, dsrc, N, iters, copies);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h, dout, 24, cudaMemcpyDeviceToHost));
    printf("rate single M=128 N=144 copies=%d: %.1f clk/MMA, %.1f clk/task, %.2f 12KB-copies per task\n", copies,
           double(h[0]) / (9.0 * iters), double(h[0]) / iters, double(h[1]) / iters);
    CK(cudaMemset(dout, 0, 64));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem2;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, rate2_probe<true>, dout, (const unsigned char*)dsrc, N, iters, copies, 16, 160));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h, dout, 24, cudaMemcpyDeviceToHost));
    printf("rate pair   M=256 N=144 copies=%d: %.1f clk/MMA, %.1f clk/task (two columns), %.2f / %.2f 12KB-copies per task in CTA 0 / 1\n",
           copies, double(h[0]) / (9.0 * iters), double(h[0]) / iters, double(h[1]) / iters, double(h[2]) / iters);
  }
  {
    const int smem3 = 6 * 2560 + 9 * 2 * 144 * 16 + 1024;
    CK(cudaFuncSetAttribute(task_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem3));
    for (int Nn : {144, 96, 48})
      for (int nc : {0, 1, 2, 3})
        for (int rot : {0, 1}) {
          if (Nn != 144 && (nc == 3 || rot == 0)) continue;
          long long h[1];
          CK(cudaMemset(dout, 0, 64));
          task_probe<<<1, 128, smem3>>>(dout, Nn, 2000, nc, rot);
          CK(cudaDeviceSynchronize());
          CK(cudaMemcpy(h, dout, 8, cudaMemcpyDeviceToHost));
          printf("task_probe N=%3d commits/task=%d rotating-slot=%d: %.1f clk/task (9 MMAs)\n", Nn, nc, rot, double(h[0]) / 2000);
        }
  }
  {
    const int smem3 = 6 * 2560 + 9 * 2 * 144 * 16 + 1024;
    CK(cudaFuncSetAttribute(dual_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem3));
    for (int Nn : {48, 64, 96})
      for (int nthr : {1, 2, 3}) {
        long long h[4] = {0, 0, 0, 0};
        CK(cudaMemset(dout, 0, 64));
        dual_probe<<<1, 128, smem3>>>(dout, Nn, 1800 / nthr, nthr);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h, dout, 32, cudaMemcpyDeviceToHost));
        long long mx = 0;
        for (int k = 0; k < nthr; ++k) mx = h[k] > mx ? h[k] : mx;
        printf("dual_probe N=%3d, %d issuing thread(s), %d MMAs in total: %.1f clk/MMA\n", Nn, nthr, 9 * (1800 / nthr) * nthr,
               double(mx) / (9.0 * (1800 / nthr) * nthr));
      }
  }
  // A-operand alignment: time taps are start-address offsets of kh * d rows (16 B each) and the K halves sit plane_rows * 16 B apart
  for (int plane_rows : {160, 117}) {
    for (int shift : {16, 8, 4, 2, 1}) {
      if (plane_rows * 16 * 5 + (2 * shift + 128) * 16 > 6 * 2560) continue;
      long long h[3];
      CK(cudaMemset(dout, 0, 64));
      rate2_probe<false><<<1, 128, smem1>>>(dout, dsrc, N, iters, 0, shift, plane_rows);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h, dout, 24, cudaMemcpyDeviceToHost));
      printf("rate single N=144, A taps shifted by %2d rows, K halves %4d B apart: %.1f clk/MMA\n", shift, plane_rows * 16,
             double(h[0]) / (9.0 * iters));
    }
  }
  return 0;
}
