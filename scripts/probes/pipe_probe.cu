// Standalone probe: issue cost (clk per warp instruction, 8 warps = 2 per scheduler) of the instruction kinds an epilogue
// can use to combine int32 accumulators, alone and WHILE one thread keeps the tensor pipe busy with tcgen05 i8 MMAs
// (N = 128, like fe3_dft_kernel).  Decides between FP64 and integer Horner steps and how to round to float32.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/probes/pipe_probe.cu -o scripts/probes/pipe_probe
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include "../../honk_b200/csrc/umma.cuh"

using namespace honk::umma;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__host__ __device__ constexpr uint32_t idesc_i8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void mma_i8(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}

enum { K_DFMA, K_DADD, K_HORNER64, K_MADWIDE, K_IMAD, K_LOP, K_F2F, K_I2F64, K_I2F32, K_FFMA, K_LL2F, K_LL2D, K_COUNT };
static const char* kNames[] = {"DFMA", "DADD", "64-bit S*256+a (C)", "mad.wide.u32 + mad.lo", "IMAD.LO", "LOP3", "F2F.F32.F64", "I2F.F64.S32",
                               "I2F.F32.S32", "FFMA", "__ll2float_rn", "__ll2double_rn"};

template <int KIND>
__global__ void __launch_bounds__(288, 1) pipe_probe(long long* out, double* sink, int iters, int with_mma, int mma_iters) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (10 * 130 * 16 + 10 * 128 * 16) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u;
  fence_async_smem();
  if (warp == 0) tmem_alloc(smem_u32(&tmem_base_s), 256);
  if (threadIdx.x == 32) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); stop = 0; }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (warp == 0) {
    if (lane == 0 && with_mma) {
      const uint32_t id = idesc_i8(128, 128);
      const long long t0 = clock64();
      long long n = 0;
      for (int r = 0; r < mma_iters && !stop; ++r) {
        for (int i = 0; i < 15; ++i) {
          const int seg = i / 5, ks = i - seg * 5;
          mma_i8(tmem + ((r & 1) ? 128 : 0), make_desc(smem_u32(smem) + (2 * ks) * 130 * 16 + seg * 16, 130 * 16, 128),
                 make_desc(smem_u32(smem + 10 * 130 * 16) + (2 * ks) * 128 * 16, 128 * 16, 128), id, i > 0);
        }
        n += 15;
        if ((r & 7) == 7) {                       // do not run unboundedly ahead of the pipe
          commit(smem_u32(&bar));
          mbar_wait(smem_u32(&bar), (r >> 3) & 1);
        }
      }
      out[2] = clock64() - t0;
      out[3] = n;
    }
  } else {
    double d0 = 1.0 + lane * 1e-9, d1 = 1.0 - lane * 1e-9, d2 = 0.5, d3 = 0.25;
    long long s0 = lane, s1 = lane * 3, s2 = lane * 5, s3 = lane * 7;
    uint32_t u0 = lane, u1 = lane + 1, u2 = lane + 2, u3 = lane + 3;
    float f0 = lane, f1 = 1.f, f2 = 2.f, f3 = 3.f;
    uint32_t v0 = lane, v1 = 1, v2 = 2, v3 = 3, w0 = lane * 77u, w1 = ~lane;
    __syncwarp();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (KIND == K_DFMA) {
          asm volatile("fma.rn.f64 %0, %0, %4, %5;\n\tfma.rn.f64 %1, %1, %4, %5;\n\tfma.rn.f64 %2, %2, %4, %5;\n\tfma.rn.f64 %3, %3, %4, %5;"
                       : "+d"(d0), "+d"(d1), "+d"(d2), "+d"(d3) : "d"(0.999), "d"(1e-3));
        } else if (KIND == K_DADD) {
          asm volatile("add.rn.f64 %0, %0, %4;\n\tadd.rn.f64 %1, %1, %4;\n\tadd.rn.f64 %2, %2, %4;\n\tadd.rn.f64 %3, %3, %4;"
                       : "+d"(d0), "+d"(d1), "+d"(d2), "+d"(d3) : "d"(1e-3));
        } else if (KIND == K_HORNER64) {
          s0 = s0 * 256 + static_cast<long long>(static_cast<int>(u0)); s1 = s1 * 256 + static_cast<long long>(static_cast<int>(u1));
          s2 = s2 * 256 + static_cast<long long>(static_cast<int>(u2)); s3 = s3 * 256 + static_cast<long long>(static_cast<int>(u3));
          asm volatile("" : "+l"(s0), "+l"(s1), "+l"(s2), "+l"(s3));
        } else if (KIND == K_MADWIDE) {
          asm volatile("{\n\t.reg .s64 a; .reg .b64 t; .reg .b32 tl, th;\n\t"
                       "cvt.s64.s32 a, %4; mad.wide.u32 t, %0, 256, a; mov.b64 {tl, th}, t; mad.lo.s32 %1, %1, 256, th; mov.b32 %0, tl;\n\t"
                       "cvt.s64.s32 a, %5; mad.wide.u32 t, %2, 256, a; mov.b64 {tl, th}, t; mad.lo.s32 %3, %3, 256, th; mov.b32 %2, tl;\n\t}"
                       : "+r"(u0), "+r"(u1), "+r"(u2), "+r"(u3) : "r"(w0), "r"(w1));
          asm volatile("{\n\t.reg .s64 a; .reg .b64 t; .reg .b32 tl, th;\n\t"
                       "cvt.s64.s32 a, %4; mad.wide.u32 t, %0, 256, a; mov.b64 {tl, th}, t; mad.lo.s32 %1, %1, 256, th; mov.b32 %0, tl;\n\t"
                       "cvt.s64.s32 a, %5; mad.wide.u32 t, %2, 256, a; mov.b64 {tl, th}, t; mad.lo.s32 %3, %3, 256, th; mov.b32 %2, tl;\n\t}"
                       : "+r"(v0), "+r"(v1), "+r"(v2), "+r"(v3) : "r"(w0), "r"(w1));
        } else if (KIND == K_IMAD) {
          asm volatile("mad.lo.s32 %0, %0, 3, %1;\n\tmad.lo.s32 %1, %1, 3, %2;\n\tmad.lo.s32 %2, %2, 3, %3;\n\tmad.lo.s32 %3, %3, 3, %0;" : "+r"(u0), "+r"(u1), "+r"(u2), "+r"(u3));
        } else if (KIND == K_LOP) {
          asm volatile("xor.b32 %0, %0, %1;\n\txor.b32 %1, %1, %2;\n\txor.b32 %2, %2, %3;\n\txor.b32 %3, %3, %0;" : "+r"(u0), "+r"(u1), "+r"(u2), "+r"(u3));
        } else if (KIND == K_F2F) {
          asm volatile("cvt.rn.f32.f64 %0, %4;\n\tcvt.rn.f32.f64 %1, %5;\n\tcvt.rn.f32.f64 %2, %6;\n\tcvt.rn.f32.f64 %3, %7;"
                       : "=f"(f0), "=f"(f1), "=f"(f2), "=f"(f3) : "d"(d0), "d"(d1), "d"(d2), "d"(d3));
        } else if (KIND == K_I2F64) {
          asm volatile("cvt.rn.f64.s32 %0, %4;\n\tcvt.rn.f64.s32 %1, %5;\n\tcvt.rn.f64.s32 %2, %6;\n\tcvt.rn.f64.s32 %3, %7;"
                       : "=d"(d0), "=d"(d1), "=d"(d2), "=d"(d3) : "r"(u0), "r"(u1), "r"(u2), "r"(u3));
        } else if (KIND == K_I2F32) {
          asm volatile("cvt.rn.f32.s32 %0, %4;\n\tcvt.rn.f32.s32 %1, %5;\n\tcvt.rn.f32.s32 %2, %6;\n\tcvt.rn.f32.s32 %3, %7;"
                       : "=f"(f0), "=f"(f1), "=f"(f2), "=f"(f3) : "r"(u0), "r"(u1), "r"(u2), "r"(u3));
        } else if (KIND == K_FFMA) {
          asm volatile("fma.rn.f32 %0, %0, %4, %5;\n\tfma.rn.f32 %1, %1, %4, %5;\n\tfma.rn.f32 %2, %2, %4, %5;\n\tfma.rn.f32 %3, %3, %4, %5;"
                       : "+f"(f0), "+f"(f1), "+f"(f2), "+f"(f3) : "f"(0.999f), "f"(1e-3f));
        } else if (KIND == K_LL2F) {
          f0 += __ll2float_rn(s0); f1 += __ll2float_rn(s1); f2 += __ll2float_rn(s2); f3 += __ll2float_rn(s3);
          s0 += 77; s1 += 77; s2 += 77; s3 += 77;
        } else if (KIND == K_LL2D) {
          d0 += __ll2double_rn(s0); d1 += __ll2double_rn(s1); d2 += __ll2double_rn(s2); d3 += __ll2double_rn(s3);
          s0 += 77; s1 += 77; s2 += 77; s3 += 77;
        }
      }
    }
    const long long t1 = clock64();
    sink[blockIdx.x * 288 + threadIdx.x] = d0 + d1 + d2 + d3 + static_cast<double>(s0 + s1 + s2 + s3) + u0 + u1 + u2 + u3 + v0 + v1 + v2 + v3 + w0 + w1 + f0 + f1 + f2 + f3;
    if (threadIdx.x == 32) out[0] = t1 - t0;
    __syncwarp();
  }
  __syncthreads();
  if (threadIdx.x == 0) stop = 1;
  fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int KIND>
void run(long long* dout, double* sink) {
  const int smem = 10 * 130 * 16 + 10 * 128 * 16 + 1024;
  CK(cudaFuncSetAttribute(pipe_probe<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 4000;
  double res[2], mma_clk = 0;
  for (int with_mma = 0; with_mma < 2; ++with_mma) {
    long long h[4] = {0, 0, 0, 0};
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaMemset(dout, 0, 64));
      pipe_probe<KIND><<<148, 288, smem>>>(dout, sink, iters, with_mma, 1 << 20);
      CK(cudaDeviceSynchronize());
    }
    CK(cudaMemcpy(h, dout, 32, cudaMemcpyDeviceToHost));
    res[with_mma] = double(h[0]) / (iters * 8.0 * 4.0);
    if (with_mma && h[3]) mma_clk = double(h[2]) / h[3];
  }
  printf("%-24s clk per group-of-one (8 warps/SM, per warp): alone %6.2f   under i8 MMA load %6.2f   (MMA %.1f clk each while probed)\n",
         kNames[KIND], res[0], res[1], mma_clk);
}

int main() {
  long long* dout;
  double* sink;
  CK(cudaMalloc(&dout, 64));
  CK(cudaMalloc(&sink, 148 * 288 * 8));
  printf("one 'instruction' = the listed operation on one value; 64-bit Horner and mad.wide are 3-instruction sequences, __ll2*_rn whatever the compiler emits\n");
  run<K_DFMA>(dout, sink); run<K_DADD>(dout, sink); run<K_HORNER64>(dout, sink); run<K_MADWIDE>(dout, sink); run<K_IMAD>(dout, sink);
  run<K_LOP>(dout, sink); run<K_F2F>(dout, sink); run<K_I2F64>(dout, sink); run<K_I2F32>(dout, sink); run<K_FFMA>(dout, sink);
  run<K_LL2F>(dout, sink); run<K_LL2D>(dout, sink);
  return 0;
}
