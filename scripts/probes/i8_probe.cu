// Standalone probe (not part of the library) for the integer-limb DFT front end:
//  (1) tcgen05.mma kind::i8 semantics with the no-swizzle K-major operand layout the conv kernels use (core matrix =
//      8 rows x 16 BYTES = 16 int8 along K, K = 32 per MMA = two chunks LBO apart), signed and unsigned A, int32
//      accumulators in TMEM, A row shifts as start-address offsets (the three hop-sized segments of a frame);
//  (2) MMA rate by N for kind::i8 vs kind::f16, one CTA and a cta_group::2 pair;
//  (3) do a DMMA (FP64 mma.sync) stream and a tcgen05.mma stream on the same SM overlap or serialise?  
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/probes/i8_probe.cu -o scripts/probes/i8_probe
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include "../../honk_b200/csrc/umma.cuh"

using namespace honk::umma;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__host__ __device__ constexpr uint32_t make_idesc_i8(int M, int N, int a_signed, int b_signed) {
  return (2u << 4) | (static_cast<uint32_t>(a_signed) << 7) | (static_cast<uint32_t>(b_signed) << 10) |
         (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void mma_i8_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma2_i8_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma2_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
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

// ---- (1) correctness: D[128][N] = A[rows shift..shift+127][K] * B[N][K]^T, K = 32 * ksteps -------------------------------
// gA: [K/16 chunks][ROWS][16 B]; gB: [K/16 chunks][N][16 B]  (the shared-memory images, copied verbatim)
__global__ void __launch_bounds__(128, 1) i8_gemm_probe(const uint8_t* __restrict__ gA, const uint8_t* __restrict__ gB, int32_t* __restrict__ gD,
                                                        int N, int ksteps, int rows, int shift, int a_signed) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int nch = 2 * ksteps;
  unsigned char* sA = smem;
  unsigned char* sB = smem + nch * rows * 16;
  for (int i = threadIdx.x; i < nch * rows * 16; i += blockDim.x) sA[i] = gA[i];
  for (int i = threadIdx.x; i < nch * N * 16; i += blockDim.x) sB[i] = gB[i];
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 256);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_i8(128, N, a_signed, 1);
    for (int ks = 0; ks < ksteps; ++ks) {
      const uint64_t ad = make_desc(smem_u32(sA) + (2 * ks) * rows * 16 + shift * 16, rows * 16, 128);
      const uint64_t bd = make_desc(smem_u32(sB) + (2 * ks) * N * 16, N * 16, 128);
      mma_i8_ss(tmem, ad, bd, idesc, ks > 0);
    }
    commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  fence_after();
  for (int c0 = 0; c0 < N; c0 += 8) {
    uint32_t v[8];
    tmem_ld8(tmem + ((warp * 32u) << 16) + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 8; ++j) gD[threadIdx.x * N + c0 + j] = static_cast<int32_t>(v[j]);
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---- (2) rate: chains of 15 accumulating MMAs (one frame tile x one twiddle digit) over a few operand tiles -----------------
// KIND 0: f16 (K = 16), 1: i8 (K = 32).  PAIR: cluster of 2 issuing cta_group::2 (B holds N/2 rows per CTA).
template <int KIND, bool PAIR>
__global__ void __launch_bounds__(128, 1) rate_probe(long long* out, int N, int iters) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  const int NB = PAIR ? N / 2 : N;
  constexpr int ROWS = 130, NCH = 10;                    // A: 10 chunks (5 MMAs) of 130 rows; B: 5 MMAs x 2 chunks x NB rows
  unsigned char* sA = smem;
  unsigned char* sB = smem + NCH * ROWS * 16;
  const int words = (NCH * ROWS * 16 + NCH * NB * 16) / 4;
  for (int i = threadIdx.x; i < words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = KIND ? 0x01010101u : 0x3c003c00u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { if (PAIR) tmem_alloc2(smem_u32(&tmem_base_s), 512); else tmem_alloc(smem_u32(&tmem_base_s), 512); }
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc = KIND ? make_idesc_i8(PAIR ? 256 : 128, N, 1, 1) : make_idesc(PAIR ? 256 : 128, N, 0);
    const long long t0 = clock64();
    for (int r = 0; r < iters; ++r) {
      const uint32_t d = tmem + ((r & 1) ? 256 : 0);     // alternate two accumulators; a chain restarts with accumulate = 0
      for (int i = 0; i < 15; ++i) {
        const int seg = i / 5, ks = i - seg * 5;
        const uint64_t ad = make_desc(smem_u32(sA) + (2 * ks) * ROWS * 16 + seg * 16, ROWS * 16, 128);
        const uint64_t bd = make_desc(smem_u32(sB) + (2 * ks) * NB * 16, NB * 16, 128);
        if (KIND) { if (PAIR) mma2_i8_ss(d, ad, bd, idesc, i > 0); else mma_i8_ss(d, ad, bd, idesc, i > 0); }
        else      { if (PAIR) mma2_f16_ss(d, ad, bd, idesc, i > 0); else mma_f16_ss(d, ad, bd, idesc, i > 0); }
      }
    }
    if (PAIR) commit2_mc(smem_u32(&bar), 3); else commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    out[0] = clock64() - t0;
  } else if (threadIdx.x == 0) {
    mbar_wait(smem_u32(&bar), 0);
  }
  fence_before();
  __syncthreads();
  if (PAIR) cluster_sync();
  if (warp == 0) { if (PAIR) tmem_dealloc2(tmem, 512); else tmem_dealloc(tmem, 512); }
}

// ---- (3) DMMA stream and tcgen05 stream on one SM: overlap or serialise? ---------------------------------------------------
__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
// mode bit 0: warp 0 issues `iters` x 9 tcgen05 f16 MMAs of N = 144; bit 1: warps 4..7 each run `diters` x 16 DMMAs
__global__ void __launch_bounds__(256, 1) mix_probe(long long* out, double* sink, int mode, int iters, int diters) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  constexpr int PLANE = 160 * 16, N = 144;
  unsigned char* sA = smem;
  unsigned char* sB = smem + 6 * PLANE;
  const int words = (6 * PLANE + 9 * 2 * N * 16) / 4;
  for (int i = threadIdx.x; i < words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  fence_async_smem();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 256);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const long long t0 = clock64();
  if (threadIdx.x == 0 && (mode & 1)) {
    const uint32_t idesc = make_idesc(128, N, 0);
    for (int r = 0; r < iters; ++r)
      for (int i = 0; i < 9; ++i) {
        const int kh = i / 3, kc = i - kh * 3;
        const uint64_t ad = make_desc(smem_u32(sA) + kc * 2 * PLANE + kh * 16, PLANE, 128);
        const uint64_t bd = make_desc(smem_u32(sB) + i * 2 * N * 16, N * 16, 128);
        mma_f16_ss(tmem, ad, bd, idesc, 1);
      }
    commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    if (blockIdx.x == 0) out[0] = clock64() - t0;
  }
  if (warp >= 4 && (mode & 2)) {
    double acc[32];
    const double a0 = 1.0 + lane * 1e-9, b0 = 1.0 - lane * 1e-9;
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = 0.0;
    for (int it = 0; it < diters; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) dmma(acc[2 * i], acc[2 * i + 1], a0, b0);
    }
    double r = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) r += acc[i];
    sink[blockIdx.x * 256 + threadIdx.x] = r;
    if (blockIdx.x == 0 && threadIdx.x == 128) out[1] = clock64() - t0;
  }
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

int main() {
  // ---- (1) correctness ----
  for (int a_signed : {1, 0})
    for (int shift : {0, 1, 2})
      for (int N : {128, 240, 64}) {
        const int ksteps = 3, K = 32 * ksteps, rows = 130, nch = 2 * ksteps;
        std::vector<uint8_t> hA(nch * rows * 16), hB(nch * N * 16);
        std::vector<int> fA(rows * K), fB(N * K);
        uint32_t s = 12345u + N + 7 * shift + a_signed;
        auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (s >> 16) & 0xff; };
        for (int r = 0; r < rows; ++r)
          for (int k = 0; k < K; ++k) {
            const uint8_t b = static_cast<uint8_t>(rnd());
            fA[r * K + k] = a_signed ? static_cast<int>(static_cast<int8_t>(b)) : static_cast<int>(b);
            hA[((k / 16) * rows + r) * 16 + k % 16] = b;
          }
        for (int n = 0; n < N; ++n)
          for (int k = 0; k < K; ++k) {
            const uint8_t b = static_cast<uint8_t>(rnd());
            fB[n * K + k] = static_cast<int>(static_cast<int8_t>(b));
            hB[((k / 16) * N + n) * 16 + k % 16] = b;
          }
        uint8_t *dA, *dB;
        int32_t* dD;
        CK(cudaMalloc(&dA, hA.size()));
        CK(cudaMalloc(&dB, hB.size()));
        CK(cudaMalloc(&dD, 128 * N * 4));
        CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xFF, 128 * N * 4));
        const int smem = nch * rows * 16 + nch * N * 16 + 1024;
        CK(cudaFuncSetAttribute(i8_gemm_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        i8_gemm_probe<<<1, 128, smem>>>(dA, dB, dD, N, ksteps, rows, shift, a_signed);
        CK(cudaDeviceSynchronize());
        std::vector<int32_t> hD(128 * N);
        CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
        long long bad = 0, maxabs = 0;
        for (int m = 0; m < 128; ++m)
          for (int n = 0; n < N; ++n) {
            long long ref = 0;
            for (int k = 0; k < K; ++k) ref += static_cast<long long>(fA[(m + shift) * K + k]) * fB[n * K + k];
            bad += ref != hD[m * N + n];
            maxabs = std::max<long long>(maxabs, ref < 0 ? -ref : ref);
          }
        printf("i8_gemm_probe A %s, row shift %d, M=128 N=%3d K=%d: %lld mismatches of %d (max |ref| %lld) %s\n", a_signed ? "s8" : "u8",
               shift, N, K, bad, 128 * N, maxabs, bad == 0 ? "EXACT" : "MISMATCH");
        cudaFree(dA); cudaFree(dB); cudaFree(dD);
      }
  // ---- (2) rate ----
  long long* dout;
  CK(cudaMalloc(&dout, 64));
  const int iters = 400;
  auto run_rate = [&](int kind, bool pair, int N) {
    const int NB = pair ? N / 2 : N;
    const int smem = 10 * 130 * 16 + 10 * NB * 16 + 1024;
    long long h[1];
    CK(cudaMemset(dout, 0, 64));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pair ? 2 : 1);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = pair ? 2 : 1;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (kind == 0 && !pair) { CK(cudaFuncSetAttribute(rate_probe<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); CK(cudaLaunchKernelEx(&cfg, rate_probe<0, false>, dout, N, iters)); }
    if (kind == 1 && !pair) { CK(cudaFuncSetAttribute(rate_probe<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); CK(cudaLaunchKernelEx(&cfg, rate_probe<1, false>, dout, N, iters)); }
    if (kind == 0 && pair) { CK(cudaFuncSetAttribute(rate_probe<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); CK(cudaLaunchKernelEx(&cfg, rate_probe<0, true>, dout, N, iters)); }
    if (kind == 1 && pair) { CK(cudaFuncSetAttribute(rate_probe<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); CK(cudaLaunchKernelEx(&cfg, rate_probe<1, true>, dout, N, iters)); }
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h, dout, 8, cudaMemcpyDeviceToHost));
    printf("rate %s %s N=%3d: %.1f clk/MMA (K=%d)  ideal N/2 = %d\n", kind ? "i8 " : "f16", pair ? "pair M=256" : "one  M=128", N,
           double(h[0]) / (15.0 * iters), kind ? 32 : 16, N / 2);
  };
  for (int kind : {0, 1})
    for (int N : {64, 96, 128, 192, 240, 256}) run_rate(kind, false, N);
  for (int kind : {0, 1})
    for (int N : {128, 192, 256}) run_rate(kind, true, N);
  // ---- (3) DMMA || tcgen05 ----
  {
    double* sink;
    CK(cudaMalloc(&sink, 148 * 256 * 8));
    const int smem = 6 * 2560 + 9 * 2 * 144 * 16 + 1024;
    CK(cudaFuncSetAttribute(mix_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int it = 3000, dit = 6000;                       // ~ 3000 * 9 * 72 = 1.9 M clk of MMAs; 6000 * 16 * 16.65 = 1.6 M clk of DMMAs per warp
    const char* names[] = {"", "tcgen05 alone", "DMMA alone (4 warps, one per scheduler)", "both"};
    for (int grid : {1, 148})
      for (int mode : {1, 2, 3}) {
        long long h[2] = {0, 0};
        for (int rep = 0; rep < 2; ++rep) {
          CK(cudaMemset(dout, 0, 64));
          mix_probe<<<grid, 256, smem>>>(dout, sink, mode, it, dit);
          CK(cudaDeviceSynchronize());
        }
        CK(cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost));
        printf("mix_probe grid=%3d %-42s: tcgen05 %.1f clk/MMA   DMMA %.2f clk/DMMA per warp\n", grid, names[mode],
               (mode & 1) ? double(h[0]) / (9.0 * it) : 0.0, (mode & 2) ? double(h[1]) / (16.0 * dit) : 0.0);
      }
  }
  return 0;
}
