// tcgen05 path for cnn-trad-pool2 (SpeechModel.forward, utils/model.py:186-205, config model.py:383):
//   conv1 1->64 k(20,8) + bias + ReLU + MaxPool(2,2)  ->  conv2 64->64 k(10,4) + bias + ReLU  ->  Linear(26624 -> L)
// One persistent CTA per SM takes whole clips; both convolutions of a clip run on the tensor cores
// back to back with the pooled conv1 output kept in shared memory.
//
// conv1 as implicit GEMM.  A pre-pass stores the MFCC image as x8[w][h] = (x[h][w], ..., x[h][w+7])
// in f16, i.e. the 8 coefficient taps of one kernel row are one 16-byte unit and time runs fastest.
// For output column w the A operand of the MMA for kernel rows (kh, kh+1) is then 128 consecutive
// units starting at row h0 + kh of column w, the second K half being the same run one row further
// (LBO = 16 bytes): M = 128 time positions (82 valid), N = 64 channels, K = 16, 10 MMAs per column,
// no im2col copy anywhere.  Output column 32 is dropped by the floor of the 2x2 pool and never computed.
// Epilogue: + bias, f16, max over the time pair (adjacent TMEM lanes, one shuffle) and the column
// pair (previous column kept in registers), ReLU, written as conv2's operand: planar
// y1[cg][w2*41 + h2][8 ch] in shared memory.  Rounding to f16 commutes with max and ReLU, so this
// equals rounding the reference's pooled FP32 value.
//
// conv2 as implicit GEMM over FLAT rows r = w*41 + h of that plane: the tap (kh, kw) of the A
// operand is the start offset kh + 41*kw, so one M = 128 tile spans about three output columns;
// rows with h >= 32 produce values that are simply not stored (78 % useful rows).  K = 64 channels
// x 40 taps = 160 MMAs per tile, 5 tiles per clip, all five accumulators live in TMEM while the
// 320 KB of conv2 weights stream once per clip through a shared-memory ring (L2-resident).
// conv2's epilogue stores relu(acc + bias) as f16 [clip][w*32 + h][64 ch]; a CUDA-core kernel does
// the Linear layer in FP32 with the weights permuted to that order at load time.
//
// Roles (10 warps): warp 0 bulk-copy producer, warp 1 MMA issuer, warps 2-5 / 6-9 two epilogue
// groups (conv1: alternate column pairs; conv2: alternate tiles).
#include "umma.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "model.cuh"

namespace honk {

using namespace umma;

namespace {

constexpr int CT_H = 101, CT_W = 40;            // MFCC frames x coefficients
constexpr int CT_C = 64;                        // feature maps of both convolutions
constexpr int CT_KH1 = 20, CT_KW1 = 8;          // conv1 kernel (time, coefficient)
constexpr int CT_HO1 = 82;                      // conv1 output rows (time)
constexpr int CT_WO1 = 32;                      // conv1 output columns that survive the pool (33 exist)
constexpr int CT_HP = 41, CT_WP = 16;           // pooled map
constexpr int CT_KH2 = 10, CT_KW2 = 4;          // conv2 kernel
constexpr int CT_HO2 = 32, CT_WO2 = 13;         // conv2 output
constexpr int CT_NPOS = CT_HO2 * CT_WO2;        // 416 positions -> 26624 features
constexpr int CT_NTILE = 5;                     // M = 128 tiles over the 13*41 = 533 flat rows
constexpr int CT_NSLAB = CT_KH2 * CT_KW2;       // weight slabs (one tap each: 4 channel chunks x 2 KB)
constexpr int CT_NS2 = 6;                       // slabs in flight
constexpr int CT_THREADS = 352;                 // warp 0 producer, 1 and 10 MMA issuers, 2-9 epilogue
constexpr int CT_X8_ROWS = CT_WO1 * CT_H;       // 3232 units per clip
constexpr int CT_X8_BYTES = CT_X8_ROWS * 16;    // 51712
constexpr int CT_X8_ALLOC = (CT_X8_ROWS + 48) * 16;                   // over-read of the last column's tile
constexpr int CT_R2 = CT_WP * CT_HP;            // 656 flat rows per channel-group plane
constexpr int CT_Y1_ALLOC = (7 * CT_R2 + 4 * 128 + 127 + (CT_KH2 - 1) + (CT_KW2 - 1) * CT_HP + 1 + 7) / 8 * 8 * 16;
constexpr int CT_W1_BYTES = (CT_KH1 / 2) * 2048;                       // 10 tiles [2 K halves][64][8] f16
constexpr int CT_SLAB_BYTES = 4 * 2048;
constexpr int CT_SMEM = CT_X8_ALLOC + CT_W1_BYTES + CT_Y1_ALLOC + CT_NS2 * CT_SLAB_BYTES;
constexpr int CT_ACT_BYTES = CT_NPOS * CT_C * 2;                       // conv2 output per clip (f16)

struct CnnTcParams {
  const unsigned char* x8;        // [n_clips][32][101][8] f16
  const unsigned char* w1;        // conv1 B tiles
  const unsigned char* w2;        // conv2 B slabs [40][4][2][64][8] f16
  const float* b1;                // [64]
  const float* b2;                // [64]
  unsigned char* act;             // [n_clips][416][64] f16
  int n_clips;
};

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
__device__ __forceinline__ uint32_t keep(uint32_t x) {
  asm volatile("mov.b32 %0, %0;" : "+r"(x));
  return x;
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ uint32_t hmax2(uint32_t a, uint32_t b) {
  uint32_t o;
  asm("max.f16x2 %0, %1, %2;" : "=r"(o) : "r"(a), "r"(b));
  return o;
}
__device__ __forceinline__ void sts16(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void stg16(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(CT_THREADS, 1) cnn_tc_kernel(const __grid_constant__ CnnTcParams p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar_xfull, bar_xfree, bar_w1;
  __shared__ __align__(8) uint64_t bar_wfull[CT_NS2], bar_wempty[CT_NS2];
  __shared__ __align__(8) uint64_t bar_c1full[3], bar_c1empty[3];
  __shared__ __align__(8) uint64_t bar_y1ready, bar_y1free, bar_c2full, bar_c2empty;
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  unsigned char* x8s = smem;
  unsigned char* w1s = x8s + CT_X8_ALLOC;
  unsigned char* y1s = w1s + CT_W1_BYTES;
  unsigned char* w2s = y1s + CT_Y1_ALLOC;
  const uint32_t a_xfull = keep(smem_u32(&bar_xfull)), a_xfree = keep(smem_u32(&bar_xfree)), a_w1 = keep(smem_u32(&bar_w1));
  const uint32_t a_wfull = keep(smem_u32(&bar_wfull[0])), a_wempty = keep(smem_u32(&bar_wempty[0]));
  const uint32_t a_c1full = keep(smem_u32(&bar_c1full[0])), a_c1empty = keep(smem_u32(&bar_c1empty[0]));
  const uint32_t a_y1ready = keep(smem_u32(&bar_y1ready)), a_y1free = keep(smem_u32(&bar_y1free));
  const uint32_t a_c2full = keep(smem_u32(&bar_c2full)), a_c2empty = keep(smem_u32(&bar_c2empty));

  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (tid == 0) {
    mbar_init(a_xfull, 1);
    mbar_init(a_xfree, 2);                      // both MMA issuers
    mbar_init(a_w1, 1);
    for (int s = 0; s < CT_NS2; ++s) {
      mbar_init(a_wfull + 8 * s, 1);
      mbar_init(a_wempty + 8 * s, 2);
    }
    for (int k = 0; k < 3; ++k) {
      mbar_init(a_c1full + 8 * k, 1);
      mbar_init(a_c1empty + 8 * k, 4);
    }
    mbar_init(a_y1ready, 8);
    mbar_init(a_y1free, 2);
    mbar_init(a_c2full, 2);
    mbar_init(a_c2empty, 8);
    fence_mbar_init();
  }
  // the tiles over-read past the rows a clip fills (tile tails, last column, last plane): keep those bytes finite
  for (int i = tid * 16; i < CT_X8_ALLOC; i += CT_THREADS * 16) *reinterpret_cast<uint4*>(x8s + i) = make_uint4(0, 0, 0, 0);
  for (int i = tid * 16; i < CT_Y1_ALLOC; i += CT_THREADS * 16) *reinterpret_cast<uint4*>(y1s + i) = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const int n_mine = (p.n_clips - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp == 0) {
    // ======================================= producer =======================================
    if (lane == 0) {
      mbar_expect_tx(a_w1, CT_W1_BYTES);
      bulk_load(smem_u32(w1s), p.w1, CT_W1_BYTES, a_w1);
      uint32_t it = 0;
      for (int k = 0; k < n_mine; ++k) {
        const long long clip = blockIdx.x + static_cast<long long>(k) * gridDim.x;
        if (k > 0) mbar_wait(a_xfree, (k - 1) & 1);                  // conv1 of the previous clip has read x8s
        mbar_expect_tx(a_xfull, CT_X8_BYTES);
        bulk_load(smem_u32(x8s), p.x8 + clip * CT_X8_BYTES, CT_X8_BYTES, a_xfull);
        for (int s = 0; s < CT_NSLAB; ++s, ++it) {
          const uint32_t st = it % CT_NS2;
          mbar_wait(a_wempty + 8 * st, ((it / CT_NS2) & 1) ^ 1);
          mbar_expect_tx(a_wfull + 8 * st, CT_SLAB_BYTES);
          bulk_load(smem_u32(w2s + st * CT_SLAB_BYTES), p.w2 + s * CT_SLAB_BYTES, CT_SLAB_BYTES, a_wfull + 8 * st);
        }
      }
    }
  } else if (warp == 1 || warp == 10) {
    // ======================================= MMA issuers ====================================
    // A tcgen05.mma costs its issuing thread ~54 clk (scripts/umma2_probe.cu: dual_probe), more than the 48 clk an N = 64
    // MMA takes to execute, so one issuing thread cannot keep the pipe full; two threads can.  No accumulator is shared:
    // in conv1 issuer 1 takes the columns that land in accumulator buffer 2 and issuer 0 those of buffers 0 and 1 (every
    // buffer's full / empty barriers then see the uses of ONE in-order issuer, which the parity waits rely on); in conv2
    // issuer 0 owns tiles 0-2 and issuer 1 tiles 3-4.
    const int q = warp == 1 ? 0 : 1;
    const int t_lo = q == 0 ? 0 : 3, t_hi = q == 0 ? 3 : CT_NTILE;
    constexpr uint32_t IDESC = make_idesc(128, CT_C, 0);
    // conv1: A K halves one row (16 B) apart, 8-row groups contiguous; B tile [2][64][8]
    const uint64_t a1_hi = make_desc(0, 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a1_lo0 = (1u << 16) | (smem_u32(x8s) >> 4);
    const uint64_t b_hi = make_desc(0, CT_C * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t b_lbo = static_cast<uint32_t>(CT_C) << 16;
    const uint32_t b1_lo0 = b_lbo | (smem_u32(w1s) >> 4);
    // conv2: A K halves one channel-group plane apart
    const uint64_t a2_hi = make_desc(0, CT_R2 * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a2_lo0 = (static_cast<uint32_t>(CT_R2) << 16) | (smem_u32(y1s) >> 4);
    const uint32_t b2_lo0 = b_lbo | (smem_u32(w2s) >> 4);
    mbar_wait(a_w1, 0);
    uint32_t cw = 0, it = 0;                    // conv1 columns and conv2 slabs issued so far
    for (int k = 0; k < n_mine; ++k) {
      mbar_wait(a_xfull, k & 1);
      fence_after();
      for (int w = 0; w < CT_WO1; ++w, ++cw) {
        const uint32_t buf = cw % 3;
        if ((buf == 2) != (q == 1)) continue;      // issuer 1 owns accumulator buffer 2, issuer 0 buffers 0 and 1
        mbar_wait(a_c1empty + 8 * buf, ((cw / 3) & 1) ^ 1);
        fence_after();
        if (elect_one()) {
          const uint32_t a_lo = a1_lo0 + w * CT_H;
          const uint32_t dcol = tmem + 5 * CT_C + buf * CT_C;
#pragma unroll
          for (int i = 0; i < CT_KH1 / 2; ++i)
            mma_f16_ss(dcol, a1_hi | static_cast<uint64_t>(a_lo + 2 * i), b_hi | static_cast<uint64_t>(b1_lo0 + i * 128), IDESC, i > 0);
          commit(a_c1full + 8 * buf);
        }
        __syncwarp();
      }
      if (elect_one()) commit(a_xfree);
      __syncwarp();
      mbar_wait(a_y1ready, k & 1);              // the pooled conv1 output of this clip is complete in y1s
      if (k > 0) mbar_wait(a_c2empty, (k - 1) & 1);   // the previous clip's conv2 accumulators have been read
      fence_after();
      for (int s = 0; s < CT_NSLAB; ++s, ++it) {
        const uint32_t st = it % CT_NS2;
        mbar_wait(a_wfull + 8 * st, (it / CT_NS2) & 1);
        fence_after();
        if (elect_one()) {
          const int kh = s / CT_KW2, kw = s - kh * CT_KW2;
          const uint32_t a_tap = a2_lo0 + kh + kw * CT_HP;
          const uint32_t b_lo = b2_lo0 + st * (CT_SLAB_BYTES / 16);
          for (int t = t_lo; t < t_hi; ++t) {
#pragma unroll
            for (int kc = 0; kc < 4; ++kc)
              mma_f16_ss(tmem + t * CT_C, a2_hi | static_cast<uint64_t>(a_tap + t * 128 + 2 * kc * CT_R2),
                         b_hi | static_cast<uint64_t>(b_lo + kc * 128), IDESC, (s > 0 || kc > 0) ? 1u : 0u);
          }
          commit(a_wempty + 8 * st);
        }
        __syncwarp();
      }
      if (elect_one()) {
        commit(a_c2full);
        commit(a_y1free);
      }
      __syncwarp();
    }
  } else {
    // ======================================= epilogue =======================================
    const int g = (warp - 2) >> 2, quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t tlane = static_cast<uint32_t>(quad * 32) << 16;
    const uint32_t y1a = smem_u32(y1s);
    for (int k = 0; k < n_mine; ++k) {
      const long long clip = blockIdx.x + static_cast<long long>(k) * gridDim.x;
      if (k > 0) mbar_wait(a_y1free, (k - 1) & 1);       // conv2 of the previous clip has read y1s
      // ---- conv1: + bias -> f16 -> max over the time pair and the column pair -> ReLU -> y1s ----
      for (int pr = g; pr < CT_WO1 / 2; pr += 2) {
        uint32_t prev[CT_C / 2];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const uint32_t cw = static_cast<uint32_t>(k) * CT_WO1 + 2 * pr + half, buf = cw % 3;
          mbar_wait(a_c1full + 8 * buf, (cw / 3) & 1);
          fence_after();
          uint32_t v[CT_C];
#pragma unroll
          for (int c = 0; c < CT_C / 16; ++c)
            tmem_ld16(tmem + tlane + 5 * CT_C + buf * CT_C + c * 16, *reinterpret_cast<uint32_t(*)[16]>(&v[c * 16]));
          tmem_ld_wait();
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(a_c1empty + 8 * buf);
#pragma unroll
          for (int j = 0; j < CT_C / 2; ++j) {
            const float2 b = __ldg(reinterpret_cast<const float2*>(p.b1) + j);
            const uint32_t t = pack_h2(__uint_as_float(v[2 * j]) + b.x, __uint_as_float(v[2 * j + 1]) + b.y);
            const uint32_t u = hmax2(t, __shfl_xor_sync(0xffffffffu, t, 1));
            if (half == 0) {
              prev[j] = u;
            } else {
              prev[j] = hmax2(hmax2(prev[j], u), 0u);
            }
          }
        }
        if ((row & 1) == 0 && row < CT_HO1) {
          const uint32_t dst = y1a + (pr * CT_HP + (row >> 1)) * 16;
#pragma unroll
          for (int cg = 0; cg < CT_C / 8; ++cg)
            sts16(dst + cg * CT_R2 * 16, prev[4 * cg], prev[4 * cg + 1], prev[4 * cg + 2], prev[4 * cg + 3]);
        }
      }
      fence_async_smem();                                 // y1s was written through the generic proxy, the MMA reads it through the async proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(a_y1ready);
      // ---- conv2: + bias -> ReLU -> f16 -> act[clip][w*32 + h][64] ----
      mbar_wait(a_c2full, k & 1);
      fence_after();
      for (int t = g; t < CT_NTILE; t += 2) {
        uint32_t v[CT_C];
#pragma unroll
        for (int c = 0; c < CT_C / 16; ++c) tmem_ld16(tmem + tlane + t * CT_C + c * 16, *reinterpret_cast<uint32_t(*)[16]>(&v[c * 16]));
        tmem_ld_wait();
        const int r = t * 128 + row;
        const int wo = r / CT_HP, ho = r - wo * CT_HP;
        if (wo < CT_WO2 && ho < CT_HO2) {
          unsigned char* dst = p.act + (clip * CT_NPOS + wo * CT_HO2 + ho) * (CT_C * 2);
#pragma unroll
          for (int cg = 0; cg < CT_C / 8; ++cg) {
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 b = __ldg(reinterpret_cast<const float2*>(p.b2) + cg * 4 + j);
              o[j] = pack_h2(fmaxf(__uint_as_float(v[cg * 8 + 2 * j]) + b.x, 0.f), fmaxf(__uint_as_float(v[cg * 8 + 2 * j + 1]) + b.y, 0.f));
            }
            stg16(dst + cg * 16, o[0], o[1], o[2], o[3]);
          }
        }
      }
      fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_c2empty);
    }
  }

  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// MFCC [n_clips][101][40] float32 -> x8[n_clips][w = 0..31][h = 0..100][8] f16 = x[h][w .. w+7]
__global__ void __launch_bounds__(256) cnn_pack_x8_kernel(const float* __restrict__ mfcc, unsigned char* __restrict__ x8) {
  __shared__ __half tile[CT_H * CT_W];
  const float* src = mfcc + static_cast<long long>(blockIdx.x) * CT_H * CT_W;
  for (int i = threadIdx.x; i < CT_H * CT_W; i += blockDim.x) tile[i] = __float2half_rn(fminf(fmaxf(__ldg(src + i), -60000.f), 60000.f));
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(x8) + static_cast<long long>(blockIdx.x) * CT_X8_ROWS;
  for (int i = threadIdx.x; i < CT_X8_ROWS; i += blockDim.x) {
    const int w = i / CT_H, h = i - w * CT_H;
    const __half* s = tile + h * CT_W + w;
    uint4 o;
    o.x = __half_as_ushort(s[0]) | (static_cast<uint32_t>(__half_as_ushort(s[1])) << 16);
    o.y = __half_as_ushort(s[2]) | (static_cast<uint32_t>(__half_as_ushort(s[3])) << 16);
    o.z = __half_as_ushort(s[4]) | (static_cast<uint32_t>(__half_as_ushort(s[5])) << 16);
    o.w = __half_as_ushort(s[6]) | (static_cast<uint32_t>(__half_as_ushort(s[7])) << 16);
    dst[i] = o;
  }
}

// Linear(26624 -> L) in FP32: logits[clip][l] = b[l] + sum_k act[clip][k] * w[l][k]  (w already in act's k order).
// One CTA = 32 clips x one K slice (blockIdx.y; split-K keeps small batches off a 104-slab serial loop; the slices'
// partial sums go to part[slice][clip][l] and cnn_fc_reduce_kernel adds them in a fixed order: deterministic);
// K in slabs of 256 staged in shared memory; thread = (clip, 32-wide k slice).
constexpr int FC_CLIPS = 32, FC_KS = 256, FC_ASTRIDE = FC_KS * 2 + 16, FC_MAXL = 16, FC_SPLIT = 13;
__global__ void __launch_bounds__(256) cnn_fc_kernel(const unsigned char* __restrict__ act, const float* __restrict__ w,
                                                     float* __restrict__ part, int n_clips, int K, int L, int k_per) {
  __shared__ __align__(16) unsigned char As[FC_CLIPS * FC_ASTRIDE];
  __shared__ __align__(16) float Ws[FC_MAXL * FC_KS];
  const int tid = threadIdx.x, c = tid & 31, ks = tid >> 5;
  const long long clip0 = static_cast<long long>(blockIdx.x) * FC_CLIPS;
  float acc[FC_MAXL];
#pragma unroll
  for (int l = 0; l < FC_MAXL; ++l) acc[l] = 0.f;
  const int k_begin = blockIdx.y * k_per, k_end = min(K, k_begin + k_per);
  for (int k0 = k_begin; k0 < k_end; k0 += FC_KS) {
    __syncthreads();
    for (int i = tid; i < FC_CLIPS * (FC_KS / 8); i += 256) {          // 16-byte units, coalesced per clip row
      const int cc = i / (FC_KS / 8), u = i - cc * (FC_KS / 8);
      uint4 v = make_uint4(0, 0, 0, 0);
      if (clip0 + cc < n_clips) v = __ldg(reinterpret_cast<const uint4*>(act + ((clip0 + cc) * K + k0) * 2) + u);
      *reinterpret_cast<uint4*>(As + cc * FC_ASTRIDE + u * 16) = v;
    }
    for (int i = tid; i < L * (FC_KS / 4); i += 256) {
      const int l = i / (FC_KS / 4), u = i - l * (FC_KS / 4);
      *reinterpret_cast<float4*>(Ws + l * FC_KS + u * 4) = __ldg(reinterpret_cast<const float4*>(w + static_cast<long long>(l) * K + k0) + u);
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; ++q) {                                       // 4 x 8 elements of this thread's slice
      const uint4 av = *reinterpret_cast<const uint4*>(As + c * FC_ASTRIDE + (ks * 32 + q * 8) * 2);
      const __half2* ah = reinterpret_cast<const __half2*>(&av);
      float a[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __half22float2(ah[j]);
        a[2 * j] = f.x;
        a[2 * j + 1] = f.y;
      }
#pragma unroll
      for (int l = 0; l < FC_MAXL; ++l) {
        if (l < L) {
          const float4 w0 = *reinterpret_cast<const float4*>(Ws + l * FC_KS + ks * 32 + q * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(Ws + l * FC_KS + ks * 32 + q * 8 + 4);
          acc[l] = fmaf(a[0], w0.x, fmaf(a[1], w0.y, fmaf(a[2], w0.z, fmaf(a[3], w0.w, acc[l]))));
          acc[l] = fmaf(a[4], w1.x, fmaf(a[5], w1.y, fmaf(a[6], w1.z, fmaf(a[7], w1.w, acc[l]))));
        }
      }
    }
  }
  __syncthreads();                                                      // As is free: reuse it for the cross-slice sums
  float(*red)[FC_CLIPS][FC_MAXL] = reinterpret_cast<float(*)[FC_CLIPS][FC_MAXL]>(As);
  static_assert(sizeof(float) * 8 * FC_CLIPS * FC_MAXL <= sizeof(As), "reduction scratch must fit in As");
#pragma unroll
  for (int l = 0; l < FC_MAXL; ++l) red[ks][c][l] = acc[l];
  __syncthreads();
  for (int i = tid; i < FC_CLIPS * L; i += 256) {
    const int cc = i / L, l = i - cc * L;
    float sum = 0.f;
    for (int s = 0; s < 8; ++s) sum += red[s][cc][l];
    if (clip0 + cc < n_clips) part[(static_cast<long long>(blockIdx.y) * n_clips + clip0 + cc) * L + l] = sum;
  }
}

__global__ void cnn_fc_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias, float* __restrict__ logits,
                                     long long n_clips, int L) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_clips * L) return;
  float sum = __ldg(bias + static_cast<int>(i % L));
  for (int s = 0; s < FC_SPLIT; ++s) sum += part[s * n_clips * L + i];
  logits[i] = sum;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct CnnTcPlan {
  int n_labels = 12, max_ctas = 148;
  unsigned char* d_w1 = nullptr;
  unsigned char* d_w2 = nullptr;
  float* d_b1 = nullptr;
  float* d_b2 = nullptr;
  float* d_wfc = nullptr;
  float* d_bfc = nullptr;
  std::vector<void*> owned;
};

void cnn_tc_plan_free(CnnTcPlan* p) {
  if (!p) return;
  for (void* q : p->owned) cudaFree(q);
  delete p;
}

namespace {
template <typename T>
int up(CnnTcPlan* p, const std::vector<T>& h, T** d) {
  void* ptr = nullptr;
  HONK_CUDA(cudaMalloc(&ptr, std::max<size_t>(h.size(), 1) * sizeof(T)));
  p->owned.push_back(ptr);
  HONK_CUDA(cudaMemcpy(ptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = static_cast<T*>(ptr);
  return HONK_OK;
}
}  // namespace

// Applies to exactly the cnn-trad-pool2 geometry (model.py:383): the kernel's tiling is specialised to it.
int cnn_tc_plan_create(honk_model* m, CnnTcPlan** out) {
  *out = nullptr;
  const honk_model_config& c = m->cfg;
  const bool fits = c.family == HONK_FAMILY_CNN && m->H0 == CT_H && m->W0 == CT_W && c.n_feature_maps1 == CT_C &&
                    c.conv1_size[0] == CT_KH1 && c.conv1_size[1] == CT_KW1 && c.conv1_stride[0] == 1 && c.conv1_stride[1] == 1 &&
                    c.conv1_pool[0] == 2 && c.conv1_pool[1] == 2 && c.has_conv2 && c.n_feature_maps2 == CT_C &&
                    c.conv2_size[0] == CT_KH2 && c.conv2_size[1] == CT_KW2 && c.conv2_stride[0] == 1 && c.conv2_stride[1] == 1 &&
                    c.conv2_pool[0] == 1 && c.conv2_pool[1] == 1 && c.tf_variant && c.dnn1_size == 0 && c.n_labels <= FC_MAXL;
  if (!fits) return HONK_OK;
  int max_smem = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
  if (CT_SMEM + 2048 > max_smem) return HONK_OK;
  std::unique_ptr<CnnTcPlan, void (*)(CnnTcPlan*)> p(new CnnTcPlan(), cnn_tc_plan_free);
  p->n_labels = c.n_labels;
  p->max_ctas = std::max(1, m->n_sms);
  {
    // conv1 B tile i (kernel rows 2i, 2i+1): [K half j][co][kw] = W[co][0][2i + j][kw]
    const HostTensor& w = m->host.at("conv1.weight");
    std::vector<__half> h(CT_W1_BYTES / 2);
    for (int i = 0; i < CT_KH1 / 2; ++i)
      for (int j = 0; j < 2; ++j)
        for (int co = 0; co < CT_C; ++co)
          for (int kw = 0; kw < CT_KW1; ++kw)
            h[((static_cast<size_t>(i) * 2 + j) * CT_C + co) * 8 + kw] = __float2half(w.data[(static_cast<size_t>(co) * CT_KH1 + 2 * i + j) * CT_KW1 + kw]);
    __half* d = nullptr;
    HONK_TRY(up(p.get(), h, &d));
    p->d_w1 = reinterpret_cast<unsigned char*>(d);
  }
  {
    // conv2 B slab (kh, kw): [kc][K half j][co][e] = W[co][kc*16 + j*8 + e][kh][kw]
    const HostTensor& w = m->host.at("conv2.weight");
    std::vector<__half> h(static_cast<size_t>(CT_NSLAB) * CT_SLAB_BYTES / 2);
    for (int kh = 0; kh < CT_KH2; ++kh)
      for (int kw = 0; kw < CT_KW2; ++kw)
        for (int kc = 0; kc < 4; ++kc)
          for (int j = 0; j < 2; ++j)
            for (int co = 0; co < CT_C; ++co)
              for (int e = 0; e < 8; ++e) {
                const int ci = kc * 16 + j * 8 + e;
                h[((((static_cast<size_t>(kh) * CT_KW2 + kw) * 4 + kc) * 2 + j) * CT_C + co) * 8 + e] =
                    __float2half(w.data[((static_cast<size_t>(co) * CT_C + ci) * CT_KH2 + kh) * CT_KW2 + kw]);
              }
    __half* d = nullptr;
    HONK_TRY(up(p.get(), h, &d));
    p->d_w2 = reinterpret_cast<unsigned char*>(d);
  }
  HONK_TRY(up(p.get(), m->host.at("conv1.bias").data, &p->d_b1));
  HONK_TRY(up(p.get(), m->host.at("conv2.bias").data, &p->d_b2));
  {
    // Linear weights in the order conv2's epilogue stores: k' = (wo*32 + ho)*64 + c  <-  flatten index c*416 + ho*13 + wo (model.py:195)
    const HostTensor& w = m->host.at("output.weight");
    const int K = CT_NPOS * CT_C;
    std::vector<float> h(static_cast<size_t>(c.n_labels) * K);
    for (int l = 0; l < c.n_labels; ++l)
      for (int wo = 0; wo < CT_WO2; ++wo)
        for (int ho = 0; ho < CT_HO2; ++ho)
          for (int ch = 0; ch < CT_C; ++ch)
            h[static_cast<size_t>(l) * K + (wo * CT_HO2 + ho) * CT_C + ch] = w.data[static_cast<size_t>(l) * K + ch * CT_NPOS + ho * CT_WO2 + wo];
    HONK_TRY(up(p.get(), h, &p->d_wfc));
    HONK_TRY(up(p.get(), m->host.at("output.bias").data, &p->d_bfc));
  }
  HONK_CUDA(cudaFuncSetAttribute(cnn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM));
  *out = p.release();
  return HONK_OK;
}

int64_t cnn_tc_workspace_bytes(const CnnTcPlan*, int64_t n_clips) {
  Arena a(nullptr, 0);
  a.take<unsigned char>(n_clips * CT_X8_BYTES);
  a.take<unsigned char>(n_clips * CT_ACT_BYTES);
  a.take<float>(n_clips * FC_SPLIT * FC_MAXL);
  return align_up(a.off, 256);
}

int cnn_tc_forward(const CnnTcPlan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                   cudaStream_t s) {
  Arena ar(ws, ws_bytes);
  unsigned char* x8 = ar.take<unsigned char>(n_clips * CT_X8_BYTES);
  unsigned char* act = ar.take<unsigned char>(n_clips * CT_ACT_BYTES);
  float* part = ar.take<float>(n_clips * FC_SPLIT * FC_MAXL);
  HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: tcgen05 workspace overrun");
  {
    ProfSpan span("cnn_pack_x8", s);
    cnn_pack_x8_kernel<<<static_cast<unsigned>(n_clips), 256, 0, s>>>(d_mfcc, x8);
    HONK_LAUNCH_CHECK();
  }
  {
    ProfSpan span("cnn_conv", s);
    CnnTcParams a{};
    a.x8 = x8;
    a.w1 = p->d_w1;
    a.w2 = p->d_w2;
    a.b1 = p->d_b1;
    a.b2 = p->d_b2;
    a.act = act;
    a.n_clips = static_cast<int>(n_clips);
    const int grid = static_cast<int>(std::min<int64_t>(p->max_ctas, n_clips));
    cnn_tc_kernel<<<grid, CT_THREADS, CT_SMEM, s>>>(a);
    HONK_LAUNCH_CHECK();
  }
  {
    ProfSpan span("cnn_fc", s);
    constexpr int K = CT_NPOS * CT_C;
    static_assert(K % (FC_SPLIT * FC_KS) == 0, "K slices must be whole slabs");
    const dim3 grid(static_cast<unsigned>((n_clips + FC_CLIPS - 1) / FC_CLIPS), FC_SPLIT);
    cnn_fc_kernel<<<grid, 256, 0, s>>>(act, p->d_wfc, part, static_cast<int>(n_clips), K, p->n_labels, K / FC_SPLIT);
    HONK_LAUNCH_CHECK();
    const long long total = n_clips * p->n_labels;
    cnn_fc_reduce_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(part, p->d_bfc, d_logits, n_clips, p->n_labels);
    HONK_LAUNCH_CHECK();
  }
  return HONK_OK;
}

}  // namespace honk
