// Standalone probe (not part of the library): sustained per-SM throughput of the two ways the
// conv kernel can stage one input column (all 48 channels, ~12 KB) into shared memory:
//   mode 0: one cp.async.bulk (1-D, contiguous 11.5 KB) per stage
//   mode 1: one 5-D cp.async.bulk.tensor box {8, 133, 1, 6, 1} (16-byte inner rows) per stage
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/tma_probe.cu -o scripts/tma_probe
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_fp16.h>
#include "../honk_b200/csrc/umma.cuh"

using namespace honk::umma;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int NSTMAX = 16;
constexpr int STAGE = 13056;

__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}

__global__ void __launch_bounds__(64, 1) probe(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap2, const unsigned char* buf,
                                               long long* out, int mode, int iters, int n_slots, int bytes, int W,
                                               long long slot_stride, long long col_stride, int NST, int split) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full[NSTMAX];
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) mbar_init(smem_u32(&full[s]), 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long g0, g1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g0));
    const long long t0 = clock64();
    for (int it = 0; it < iters + NST; ++it) {
      const int s = it % NST;
      if (it >= NST) mbar_wait(smem_u32(&full[s]), ((it / NST) - 1) & 1);
      if (it < iters) {
        const int task = blockIdx.x * iters + it;
        const int slot = (task / W) % n_slots, w = task % W;
        const uint32_t dst = smem_u32(smem + s * STAGE);
        const uint32_t fb = smem_u32(&full[s]);
        mbar_expect_tx(fb, bytes);
        if (mode == 0) {
          const int part = bytes / split;
          for (int p = 0; p < split; ++p) bulk_load(dst + p * part, buf + slot * slot_stride + w * col_stride + p * part, part, fb);
        }
        else if (mode == 1) tma_load_5d(dst, &tmap, fb, 0, -16, w, 0, slot);
        else tma_load_5d(dst, &tmap2, fb, 0, 0, 0, w, slot);
      }
    }
    const long long t1 = clock64();
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
    out[blockIdx.x] = t1 - t0;
    out[148 + blockIdx.x] = (long long)(g1 - g0);
  }
}

int main() {
  const int H = 101, W = 40, NCG = 6;
  PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  void* fn = nullptr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  long long* dT;
  CK(cudaMalloc(&dT, 2 * 148 * 8));
  for (int n_slots : {74}) {     // 74 slots: L2 resident; 2048 slots: streams from HBM
    // mode 0 layout: [slot][w][cg][117 rows][8] halves (+ tail)
    const long long col_stride = (long long)NCG * 117 * 16, slot_stride = col_stride * W + 4096;
    // mode 1 layout: [slot][cg][w][h][8]
    const size_t bytes_total = (size_t)n_slots * (11520ll * W + 4096) + 65536;
    unsigned char* buf;
    CK(cudaMalloc(&buf, bytes_total));
    CK(cudaMemset(buf, 0, bytes_total));
    CUtensorMap tm;
    const cuuint64_t dims[5] = {8, (cuuint64_t)H, (cuuint64_t)W, (cuuint64_t)NCG, (cuuint64_t)n_slots};
    const cuuint64_t strides[4] = {16, 16ull * H, 16ull * H * W, 16ull * H * W * NCG};
    cuuint32_t box[5] = {8, 133, 1, 6, 1};
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    CUtensorMap tm2;
    {
      const cuuint64_t dims2[5] = {64, 17, (cuuint64_t)NCG, (cuuint64_t)W, (cuuint64_t)n_slots};
      const cuuint64_t strides2[4] = {128, 1872, 11232, 11232ull * W + 4096};
      cuuint32_t box2[5] = {64, 17, 6, 1, 1};
      r = encode(&tm2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, buf, dims2, strides2, box2, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("encode2 failed %d\n", (int)r); return 1; }
    }
        CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, NSTMAX * STAGE));
    struct Cfg { int mode, bytes, nst, split; long long cs; };
    const Cfg cfgs[] = {{0, 11488, 8, 1, 11232}, {0, 11520, 4, 1, 11520}, {0, 11520, 8, 1, 11520},
                        {0, 11520, 8, 6, 11520},
                        {1, 12768, 4, 1, 0}, {2, 13056, 2, 1, 0}, {2, 13056, 4, 1, 0}, {2, 13056, 8, 1, 0}};
    for (int grid : {1, 148}) for (const Cfg& c : cfgs) {
      const int iters = 2000;
      const long long cs = c.cs ? c.cs : col_stride;
      for (int rep = 0; rep < 2; ++rep) {
        probe<<<grid, 64, NSTMAX * STAGE>>>(tm, tm2, buf, dT, c.mode, iters, n_slots, c.bytes, W, cs * W + 4096, cs, c.nst, c.split);
        CK(cudaDeviceSynchronize());
      }
      std::vector<long long> hT(grid);
      CK(cudaMemcpy(hT.data(), dT, grid * 8, cudaMemcpyDeviceToHost));
      long long mx = 0;
      for (auto t : hT) mx = t > mx ? t : mx;
      long long ns; CK(cudaMemcpy(&ns, dT + 148, 8, cudaMemcpyDeviceToHost));
      printf("tma_probe slots=%4d grid=%3d mode=%d nst=%2d split=%2d: %5d B/stage, %.0f clk/stage, %.1f B/clk/SM, %.0f ns/stage (%.2f GHz)\n", n_slots, grid, c.mode,
             c.nst, c.split, c.bytes, (double)mx / iters, (double)c.bytes * iters / mx, (double)ns / iters, (double)mx / ns);
    }
    CK(cudaFree(buf));
  }
  return 0;
}
