// Dilated 3x3 convolution C->C of the res* models as a TMA-fed tcgen05 implicit GEMM with the
// layer's elementwise tail fused into the epilogue (SpeechResModel.forward, utils/model.py:107-118):
//     y = relu(conv_i(x));  x = y (+ old_x, old_x = x when i is even);  x = bn_i(x)
//
// GEMM view per unit: M = 128 consecutive positions of the packed input columns (time fastest),
// N = C padded (48 / 32), K = 16 channels per tcgen05.mma, 9 taps x NKC chunks accumulate into one
// TMEM tile.  A tap is a start-address offset into the same shared-memory stage; dilation in the
// coefficient axis is removed by loading every d-th column (strided TMA box), dilation in time is
// an offset of d rows, zero padding comes from TMA out-of-bounds fill.
//
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (elect.sync, precomputed descriptors),
// warps 2-5 = epilogue (TMEM -> registers -> ReLU / residual / BN -> f16 -> HBM).
#include "res_tc.cuh"
#include "umma.cuh"

namespace honk {

using namespace umma;

namespace {

constexpr int MAX_STAGES = 6;
constexpr int SMEM_FRONT_PAD = 1024;

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  a = fminf(fmaxf(a, -65504.f), 65504.f);
  b = fminf(fmaxf(b, -65504.f), 65504.f);
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

template <int NCG>
__global__ void __launch_bounds__(TC_THREADS, 1)
tc_conv_kernel(const __grid_constant__ CUtensorMap tmap, const TcConvArgs a) {
  constexpr int N = NCG * 8;            // GEMM N (padded output channels)
  constexpr int NKC = NCG / 2;          // K chunks of 16 input channels
  constexpr int NSLOT = 512 / N;        // TMEM accumulator tiles (10 for N=48, 16 for N=32)
  constexpr int WBYTES = 9 * NCG * N * 16;

  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar_full[MAX_STAGES], bar_empty[MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_acc_full[16], bar_acc_empty[16];
  __shared__ uint32_t tmem_base_s;
  __shared__ float s_scale[N], s_shift[N];

  const TcLayerGeom& g = a.g;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* stages = smem + SMEM_FRONT_PAD;
  unsigned char* wsm = stages + g.n_stages * g.stage_stride;

  // ---- one-time setup -----------------------------------------------------------------------
  {
    const uint4* src = reinterpret_cast<const uint4*>(a.wpack);
    uint4* dst = reinterpret_cast<uint4*>(wsm);
    for (int i = tid; i < WBYTES / 16; i += TC_THREADS) dst[i] = __ldg(src + i);
    // zero the pads: taps of the last rows read d rows past a column / stage
    for (int i = tid; i < SMEM_FRONT_PAD / 16; i += TC_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    const int pad16 = (g.stage_stride - g.stage_bytes) / 16;
    for (int s = 0; s < g.n_stages; ++s)
      for (int i = tid; i < pad16; i += TC_THREADS)
        reinterpret_cast<uint4*>(stages + s * g.stage_stride + g.stage_bytes)[i] = make_uint4(0, 0, 0, 0);
    if (tid < N) {
      s_scale[tid] = a.bn_scale[tid];
      s_shift[tid] = a.bn_shift[tid];
    }
    fence_async_smem();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (tid == 0) {
    for (int s = 0; s < g.n_stages; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int k = 0; k < NSLOT; ++k) {
      mbar_init(smem_u32(&bar_acc_full[k]), 1);
      mbar_init(smem_u32(&bar_acc_empty[k]), 4);
    }
    fence_mbar_init();
    prefetch_tmap(&tmap);
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const int total_units = a.n_clips * g.n_units;

  if (warp == 0) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      int it = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int clip = u / g.n_units;
        const TcUnit un = g.units[u - clip * g.n_units];
        for (int kc = 0; kc < NKC; ++kc, ++it) {
          const int s = it % g.n_stages;
          const uint32_t ph = (it / g.n_stages) & 1;
          mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&bar_full[s]);
          mbar_expect_tx(full, g.stage_bytes);
          const uint32_t dst = smem_u32(stages + s * g.stage_stride);
          if (g.strided) {
            tma_load_5d(dst, &tmap, full, 0, -g.d, un.w_first, 2 * kc, clip);
          } else {
            for (int p = 0; p < 2; ++p)
              for (int c = 0; c < g.box_cols; ++c)
                tma_load_5d(dst + (p * g.box_cols + c) * g.Hp * 16, &tmap, full, 0, -g.d, un.w_first + c * g.d,
                            2 * kc + p, clip);
          }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ================================
    constexpr uint32_t idesc = make_idesc(128, N, 0);
    const uint32_t plane16 = static_cast<uint32_t>(g.box_cols * g.Hp);       // A plane stride, 16-B units
    const uint64_t a_hi = make_desc(0, plane16 * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a_lbo = (plane16 & 0x3FFF) << 16;
    const uint64_t b_hi = make_desc(0, N * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t b_lo0 = ((static_cast<uint32_t>(N) & 0x3FFF) << 16) | (smem_u32(wsm) >> 4);
    int tap_off[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) tap_off[t] = (t % 3) * g.Hp + (t / 3 - 1) * g.d;   // kw -> column, kh -> time
    int it = 0, gt = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int clip = u / g.n_units;
      const int T = g.units[u - clip * g.n_units].n_tiles;
      for (int kc = 0; kc < NKC; ++kc, ++it) {
        const int s = it % g.n_stages;
        mbar_wait(smem_u32(&bar_full[s]), (it / g.n_stages) & 1);
        fence_after();
        const uint32_t a_lo0 = a_lbo | (smem_u32(stages + s * g.stage_stride) >> 4);
        for (int m = 0; m < T; ++m) {
          const int gm = gt + m, slot = gm % NSLOT;
          if (kc == 0) {
            mbar_wait(smem_u32(&bar_acc_empty[slot]), ((gm / NSLOT) & 1) ^ 1);
            fence_after();
          }
          if (elect_one()) {
            const uint32_t acc = tmem + slot * N;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const uint64_t ad = a_hi | static_cast<uint64_t>(a_lo0 + static_cast<uint32_t>(m * 128 + tap_off[t]));
              const uint64_t bd = b_hi | static_cast<uint64_t>(b_lo0 + static_cast<uint32_t>((t * NCG + 2 * kc) * N));
              mma_f16_ss(acc, ad, bd, idesc, (kc | t) != 0);
            }
            if (kc == NKC - 1) commit(smem_u32(&bar_acc_full[slot]));
          }
          __syncwarp();
        }
        if (elect_one()) commit(smem_u32(&bar_empty[s]));
        __syncwarp();
      }
      gt += T;
    }
  } else {
    // =============================== epilogue ==================================
    const int quad = warp & 3;                 // TMEM lanes 32*quad .. 32*quad+31
    const bool add_res = a.flags & 1, last = a.flags & 2;
    const long long plane_elems = static_cast<long long>(g.W) * g.H * 8;
    int gt = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int clip = u / g.n_units;
      const TcUnit un = g.units[u - clip * g.n_units];
      for (int m = 0; m < un.n_tiles; ++m, ++gt) {
        const int slot = gt % NSLOT;
        const int q = m * 128 + quad * 32 + lane;
        const int c = q / g.Hp, r = q - c * g.Hp;
        const int w = un.w_first + (c + 1) * g.d, h = r - g.d;
        const bool valid = (h >= 0) && (h < g.H) && (c < un.n_cols);   // Hp may exceed H + d (alignment rows)
        const long long base = (static_cast<long long>(clip) * NCG * g.W * g.H + static_cast<long long>(w) * g.H + h) * 8;
        uint4 rv[NCG];
        if (add_res && valid) {
#pragma unroll
          for (int cg = 0; cg < NCG; ++cg) rv[cg] = *reinterpret_cast<const uint4*>(a.res + base + cg * plane_elems);
        }
        mbar_wait(smem_u32(&bar_acc_full[slot]), (gt / NSLOT) & 1);
        fence_after();
        uint32_t v[N];
        const uint32_t taddr = tmem + (static_cast<uint32_t>(quad * 32) << 16) + slot * N;
#pragma unroll
        for (int j = 0; j < N / 16; ++j) tmem_ld16(taddr + j * 16, *reinterpret_cast<uint32_t(*)[16]>(&v[j * 16]));
        tmem_ld_wait();
        fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_acc_empty[slot]));

#pragma unroll
        for (int cg = 0; cg < NCG; ++cg) {
          float x[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) x[k] = fmaxf(__uint_as_float(v[cg * 8 + k]), 0.f);
          if (add_res && valid) {
            const __half2* rh = reinterpret_cast<const __half2*>(&rv[cg]);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float2 f = __half22float2(rh[k]);
              x[2 * k] += f.x;
              x[2 * k + 1] += f.y;
            }
            uint4 o;
            o.x = pack_h2(x[0], x[1]); o.y = pack_h2(x[2], x[3]); o.z = pack_h2(x[4], x[5]); o.w = pack_h2(x[6], x[7]);
            *reinterpret_cast<uint4*>(a.res + base + cg * plane_elems) = o;
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) x[k] = fmaf(x[k], s_scale[cg * 8 + k], s_shift[cg * 8 + k]);
          if (!last) {
            if (valid) {
              uint4 o;
              o.x = pack_h2(x[0], x[1]); o.y = pack_h2(x[2], x[3]); o.z = pack_h2(x[4], x[5]); o.w = pack_h2(x[6], x[7]);
              *reinterpret_cast<uint4*>(a.out + base + cg * plane_elems) = o;
            }
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              float sum = valid ? x[k] : 0.f;
#pragma unroll
              for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
              if (lane == 0) atomicAdd(a.pool + static_cast<long long>(clip) * N + cg * 8 + k, sum);
            }
          }
        }
      }
    }
  }

  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace

int tc_conv_prepare(int ncg, int max_smem) {
  if (ncg == 6) HONK_CUDA(cudaFuncSetAttribute(tc_conv_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  else if (ncg == 4) HONK_CUDA(cudaFuncSetAttribute(tc_conv_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  else return set_error(HONK_ERR_UNSUPPORTED, "tcgen05 conv: unsupported channel groups %d", ncg);
  return HONK_OK;
}

int tc_conv_launch(int ncg, const TcConvArgs& a, const CUtensorMap& tmap, int grid, cudaStream_t s) {
  if (ncg == 6) tc_conv_kernel<6><<<grid, TC_THREADS, a.g.smem_bytes, s>>>(tmap, a);
  else tc_conv_kernel<4><<<grid, TC_THREADS, a.g.smem_bytes, s>>>(tmap, a);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

}  // namespace honk
