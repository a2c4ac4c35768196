// Probe 2: burst of n TMA loads issued back to back, then waited: T(n) = latency + n * bytes / rate.
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
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__global__ void __launch_bounds__(64, 1) probe(const __grid_constant__ CUtensorMap tmap2, const unsigned char* buf, long long* out,
                                               int mode, int n, int bytes, int reps) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t full[NSTMAX];
  if (threadIdx.x == 0) { for (int s = 0; s < NSTMAX; ++s) mbar_init(smem_u32(&full[s]), 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long best = 1ll << 60, issue = 0;
    for (int rep = 0; rep < reps; ++rep) {
      const long long t0 = clock64();
      for (int s = 0; s < n; ++s) {
        const uint32_t fb = smem_u32(&full[s]);
        mbar_expect_tx(fb, bytes);
        const int w = (rep * n + s + blockIdx.x * 7) % 40, slot = (rep + blockIdx.x) % 64;
        if (mode == 0) bulk_load(smem_u32(smem + s * STAGE), buf + slot * (11232ll * 40 + 4096) + w * 11232, bytes, fb);
        else tma_load_5d(smem_u32(smem + s * STAGE), &tmap2, fb, 0, 0, 0, w, slot);
      }
      const long long t1 = clock64();
      for (int s = 0; s < n; ++s) mbar_wait(smem_u32(&full[s]), rep & 1);
      const long long t2 = clock64();
      if (t2 - t0 < best) { best = t2 - t0; issue = t1 - t0; }
    }
    out[blockIdx.x] = best;
    out[148 + blockIdx.x] = issue;
  }
}
int main() {
  const int W = 40, NCG = 6, n_slots = 74;
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  auto encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  long long* dT; CK(cudaMalloc(&dT, 2 * 148 * 8));
  unsigned char* buf; size_t total = (size_t)n_slots * (11520ll * W + 4096) + 65536;
  CK(cudaMalloc(&buf, total)); CK(cudaMemset(buf, 0, total));
  CUtensorMap tm2;
  const cuuint64_t dims2[5] = {64, 17, (cuuint64_t)NCG, (cuuint64_t)W, (cuuint64_t)n_slots};
  const cuuint64_t strides2[4] = {128, 1872, 11232, 11232ull * W + 4096};
  cuuint32_t box2[5] = {64, 17, 6, 1, 1}; cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = encode(&tm2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, buf, dims2, strides2, box2, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode2 failed %d\n", (int)r); return 1; }
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, NSTMAX * STAGE));
  for (int grid : {1, 148}) for (int mode : {0, 2}) for (int bytes : {1024, 4096, 11520}) for (int n : {1, 2, 4, 8, 16}) {
    if (mode == 2 && bytes != 11520) continue;
    const int b = mode == 2 ? 13056 : bytes;
    probe<<<grid, 64, NSTMAX * STAGE>>>(tm2, buf, dT, mode, n, b, 20);
    CK(cudaDeviceSynchronize());
    std::vector<long long> hT(2 * 148);
    CK(cudaMemcpy(hT.data(), dT, 2 * 148 * 8, cudaMemcpyDeviceToHost));
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = hT[i] > mx ? hT[i] : mx;
    printf("burst grid=%3d mode=%d bytes=%5d n=%2d: total %6lld clk (issue %5lld) -> %.1f B/clk\n", grid, mode, b, n, mx, hT[148], (double)b * n / mx);
  }
  return 0;
}
