// Persistent multi-layer tcgen05 kernel for the un-pooled res* models (res15, res15-narrow):
// every dilated 3x3 C->C layer of SpeechResModel.forward (utils/model.py:104-121) plus conv0 in
// ONE launch, with the activations of the clips in flight kept L2-resident.
//
// GEMM formulation ("ring accumulate").  One task = one INPUT column (MFCC-coefficient index w)
// of one clip: M = 128 time positions, K = all input channels, and the three coefficient taps kw
// are stacked into N = 3*NB (144 for 45->48 channels), so one tcgen05.mma is MMA-bound instead of
// shared-memory-operand bound (N = 48 reaches only 54 % of the pipe, scripts/umma_probe.cu).
// Input column c contributes its kw block to OUTPUT column c - (kw-1)*d.  Columns of one residue
// class mod d form a chain; along a chain the three N blocks of input j land on outputs j-1, j,
// j+1, which sit in ADJACENT 48-column blocks of a TMEM ring, so the 3-way shifted add happens
// inside the tensor core accumulator (no cross-lane traffic, no extra TMEM reads).  Time taps kh
// are start-address offsets of the A operand (rows of the staged column), time dilation = kh*d
// rows, zero padding = physically zero rows between the channel planes in HBM.
//
// BatchNorm (affine=False, after the residual add, model.py:112-118) is folded exactly: the
// stored activation is z_i = x_i - mean_i, so zero padding of z equals the reference's zero
// padding of bn_i(x_i); the scale 1/sqrt(var_i+eps) is folded into the NEXT layer's weights, and
// the raw residual is recovered as x_{i-2} = z_{i-2} + mean_{i-2} from the buffer the previous
// layer read.  Even layers therefore update their residual buffer in place and two rotating
// buffers (P, Q) per clip suffice.
//
// Scheduling.  A team of 4 CTAs (consecutive blockIdx) owns two clips at a time and runs them
// through all layers, alternating clips per layer so that the team-wide "layer done" flag of one
// clip is hidden behind the other clip's work.  Each CTA owns whole chains (or chain segments
// with one halo input column per cut).  Roles: warp 0 = bulk-copy (TMA) producer, warp 1 = MMA
// issuer, warps 2-9 = epilogue (TMEM -> ReLU + residual + mean shift -> f16 -> L2) and conv0.
#include "res_tc.cuh"
#include "umma.cuh"

#include <algorithm>
#include <cmath>
#include <vector>

#include "model.cuh"

namespace honk {

using namespace umma;

constexpr int T2_G = 4;           // CTAs per team
constexpr int T2_NSTAGE = 8;      // A-operand stages (one input column each)
constexpr int T2_THREADS = 320;
constexpr int T2_PAD = 16;        // zero rows above every channel plane (>= max dilation)
constexpr int T2_MAXSEG = 8;
constexpr int T2_EPI_THREADS = 256;

struct T2Seg {                    // a run of output columns of one chain, owned by one CTA
  short r;                        // residue class (first column of the chain)
  short j0, n_out;                // chain indices of the outputs
  short jin0, n_in;               // chain indices of the inputs (outputs +- 1, clamped to the chain)
  short pad[3];
};

struct T2Layer {
  int d;                          // dilation
  int flags;                      // bit0: add residual, bit1: last layer (pool, no store)
  int w_off;                      // byte offset of this layer's packed weights
  int cst_off;                    // float offset of the epilogue constants [NB]
  int n_seg[T2_G];
  T2Seg seg[T2_G][T2_MAXSEG];
};

struct T2Params {
  const T2Layer* layers;
  const unsigned char* wpack;
  const float* cst;
  const float* mfcc;              // [n_clips][H][W] float32
  const float* w0;                // conv0 weights [9][NB]
  unsigned char* P;               // [n_slots] activation buffers
  unsigned char* Q;               // [2][n_slots]
  float* pool;                    // [n_clips][NB]
  unsigned int* flags;            // [n_clips] team-wide layer counters
  long long slot_stride;          // bytes per clip buffer
  int n_layers, n_clips, n_slots;
  int H, W;
  int plane_rows;                 // H + T2_PAD
  int col_bytes;                  // NCG * plane_rows * 16
  int stage_stride, wbytes;
};

namespace {

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void wait_flag(const unsigned int* p, unsigned int target) {
  for (uint32_t i = 0; ld_acquire_gpu(p) < target; ++i) {
    __nanosleep(40);
    if (i > (1u << 25)) __trap();
  }
}
__device__ __forceinline__ uint4 ldcg16(const void* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void stcg16(void* p, const uint4& v) {
  asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  a = fminf(fmaxf(a, -65504.f), 65504.f);
  b = fminf(fmaxf(b, -65504.f), 65504.f);
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

template <int NCG>
__global__ void __launch_bounds__(T2_THREADS, 1) tc2_res_kernel(const T2Params p) {
  constexpr int NB = NCG * 8;            // padded channels = N block
  constexpr int NKC = NCG / 2;           // K chunks of 16 channels
  constexpr int NSLOT = 512 / NB;        // TMEM ring of output-column accumulators
  constexpr int HALF = NB / 2;           // columns per epilogue warp
  constexpr int BT16 = 6 * NB;           // one (kh,kc) B tile in 16-byte units: 2 K-halves x 3NB rows
  constexpr int NMMA = 3 * NKC;          // MMAs per input column

  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar_full[T2_NSTAGE], bar_empty[T2_NSTAGE];
  __shared__ __align__(8) uint64_t bar_acc_full[NSLOT], bar_acc_empty[NSLOT];
  __shared__ __align__(8) uint64_t bar_wfull[2], bar_wfree[2];
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int team = blockIdx.x / T2_G, rank = blockIdx.x % T2_G, n_teams = gridDim.x / T2_G;
  unsigned char* stages = smem;
  unsigned char* wbuf = smem + T2_NSTAGE * p.stage_stride;
  float* xs = reinterpret_cast<float*>(wbuf + 2 * p.wbytes);          // conv0 input tile [(H+2)][XS]
  const int XS = p.W + 3;
  float* w0s = xs + (p.H + 2) * XS;                                   // [9][NB]

  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (tid == 0) {
    for (int s = 0; s < T2_NSTAGE; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int k = 0; k < NSLOT; ++k) {
      mbar_init(smem_u32(&bar_acc_full[k]), 1);
      mbar_init(smem_u32(&bar_acc_empty[k]), 8);
    }
    for (int k = 0; k < 2; ++k) {
      mbar_init(smem_u32(&bar_wfull[k]), 1);
      mbar_init(smem_u32(&bar_wfree[k]), 1);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < 9 * NB; i += T2_THREADS) w0s[i] = __ldg(p.w0 + i);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const int n_pairs = (p.n_clips + 1) / 2;
  const int plane_bytes = p.plane_rows * 16;

  if (warp == 0) {
    // ======================================= producer =======================================
    if (lane == 0) {
      uint32_t it = 0, q = 0;
      int piter = 0;
      for (int pair = team; pair < n_pairs; pair += n_teams, ++piter) {
        const int nab = min(2, p.n_clips - 2 * pair);
        for (int L = 1; L <= p.n_layers; ++L, ++q) {
          const T2Layer* ly = p.layers + (L - 1);
          if (q == 0) {
            mbar_expect_tx(smem_u32(&bar_wfull[0]), p.wbytes);
            bulk_load(smem_u32(wbuf), p.wpack + __ldg(&ly->w_off), p.wbytes, smem_u32(&bar_wfull[0]));
          }
          const int d = __ldg(&ly->d);
          const uint32_t bytes = (NCG - 1) * plane_bytes + (p.H + 2 * d) * 16;
          const int nseg = __ldg(&ly->n_seg[rank]);
          for (int ab = 0; ab < nab; ++ab) {
            const int clip = 2 * pair + ab, slot = 2 * team + ab;
            const unsigned char* in = (L & 1) ? p.Q + (static_cast<long long>(piter & 1) * p.n_slots + slot) * p.slot_stride
                                              : p.P + static_cast<long long>(slot) * p.slot_stride;
            in += (T2_PAD - d) * 16;
            wait_flag(p.flags + clip, T2_G * L);          // every CTA of the team finished layer L-1 (conv0 = layer 0)
            fence_proxy_async_global();
            for (int s = 0; s < nseg; ++s) {
              const T2Seg* sg = &ly->seg[rank][s];
              const int r = __ldg(&sg->r), jin0 = __ldg(&sg->jin0), n_in = __ldg(&sg->n_in);
              for (int t = 0; t < n_in; ++t, ++it) {
                const int st = it % T2_NSTAGE;
                mbar_wait(smem_u32(&bar_empty[st]), ((it / T2_NSTAGE) & 1) ^ 1);
                const uint32_t full = smem_u32(&bar_full[st]);
                mbar_expect_tx(full, bytes);
                bulk_load(smem_u32(stages + st * p.stage_stride), in + static_cast<long long>(r + (jin0 + t) * d) * p.col_bytes,
                          bytes, full);
              }
            }
          }
          // prefetch the weights of the next layer in sequence into the other buffer
          const bool more = (L < p.n_layers) || (pair + n_teams < n_pairs);
          if (more) {
            const uint32_t qn = q + 1;
            const T2Layer* nl = p.layers + (L < p.n_layers ? L : 0);
            if (qn >= 2) mbar_wait(smem_u32(&bar_wfree[qn & 1]), ((qn - 2) >> 1) & 1);
            mbar_expect_tx(smem_u32(&bar_wfull[qn & 1]), p.wbytes);
            bulk_load(smem_u32(wbuf + (qn & 1) * p.wbytes), p.wpack + __ldg(&nl->w_off), p.wbytes, smem_u32(&bar_wfull[qn & 1]));
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer =====================================
    const uint64_t a_hi = make_desc(0, plane_bytes, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a_lbo = (static_cast<uint32_t>(p.plane_rows) & 0x3FFF) << 16;
    const uint64_t b_hi = make_desc(0, 3 * NB * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t b_lbo = (static_cast<uint32_t>(3 * NB) & 0x3FFF) << 16;
    uint32_t it = 0, q = 0, g = 0;
    for (int pair = team; pair < n_pairs; pair += n_teams) {
      const int nab = min(2, p.n_clips - 2 * pair);
      for (int L = 1; L <= p.n_layers; ++L, ++q) {
        const T2Layer* ly = p.layers + (L - 1);
        const int d = __ldg(&ly->d);
        const int nseg = __ldg(&ly->n_seg[rank]);
        mbar_wait(smem_u32(&bar_wfull[q & 1]), (q >> 1) & 1);
        const uint32_t b_lo0 = b_lbo | (smem_u32(wbuf + (q & 1) * p.wbytes) >> 4);
        for (int ab = 0; ab < nab; ++ab) {
          for (int s = 0; s < nseg; ++s) {
            const T2Seg* sg = &ly->seg[rank][s];
            const int j0 = __ldg(&sg->j0), n_out = __ldg(&sg->n_out), jin0 = __ldg(&sg->jin0), n_in = __ldg(&sg->n_in);
            const int jend = j0 + n_out - 1;
            for (int t = 0; t < n_in; ++t, ++it) {
              const int jc = jin0 + t;
              const int st = it % T2_NSTAGE;
              const int o_lo = max(jc - 1, j0), o_hi = min(jc + 1, jend);
              const int ft_lo = (t == 0) ? o_lo : jc + 1;                  // first-touched outputs: ft_lo..o_hi
              for (int o = ft_lo; o <= o_hi; ++o) {
                const uint32_t gi = g + (o - j0);
                mbar_wait(smem_u32(&bar_acc_empty[gi % NSLOT]), ((gi / NSLOT) & 1) ^ 1);
              }
              mbar_wait(smem_u32(&bar_full[st]), (it / T2_NSTAGE) & 1);
              fence_after();
              if (elect_one()) {
                const uint32_t a_lo0 = a_lbo | (smem_u32(stages + st * p.stage_stride) >> 4);
                // issues D[outs lo..hi] (+)= A(kh,kc) * B(kh,kc)[blocks], split where the ring wraps
                auto issue = [&](int lo, int hi, int i, uint32_t acc) {
                  const int n = hi - lo + 1;
                  if (n <= 0) return;
                  const int kh = i / NKC, kc = i - kh * NKC;
                  const uint64_t ad = a_hi | static_cast<uint64_t>(a_lo0 + static_cast<uint32_t>(2 * kc * p.plane_rows + kh * d));
                  const uint32_t gi = g + (lo - j0);
                  const int s_lo = gi % NSLOT;
                  const int first = min(n, NSLOT - s_lo);
                  const int blk = lo - (jc - 1);
                  const uint64_t bd = b_hi | static_cast<uint64_t>(b_lo0 + static_cast<uint32_t>(i * BT16 + blk * NB));
                  mma_f16_ss(tmem + s_lo * NB, ad, bd, make_idesc(128, first * NB, 0), acc);
                  if (n > first)
                    mma_f16_ss(tmem, ad, bd + static_cast<uint64_t>(first * NB), make_idesc(128, (n - first) * NB, 0), acc);
                };
                issue(o_lo, min(ft_lo - 1, o_hi), 0, 1);
                issue(max(ft_lo, o_lo), o_hi, 0, 0);
#pragma unroll
                for (int i = 1; i < NMMA; ++i) issue(o_lo, o_hi, i, 1);
                commit(smem_u32(&bar_empty[st]));
                if (jc - 1 >= j0 && jc - 1 <= jend) commit(smem_u32(&bar_acc_full[(g + (jc - 1 - j0)) % NSLOT]));
                if (t == n_in - 1 && jc <= jend) commit(smem_u32(&bar_acc_full[(g + (jc - j0)) % NSLOT]));
              }
              __syncwarp();
            }
            g += n_out;
          }
        }
        if (elect_one()) commit(smem_u32(&bar_wfree[q & 1]));
        __syncwarp();
      }
    }
  } else {
    // ======================================= epilogue =======================================
    const int ew = warp - 2, quad = warp & 3, half = ew >> 2;
    const int h = quad * 32 + lane;
    const bool valid = h < p.H;
    const int etid = ew * 32 + lane;
    uint32_t g = 0;
    int piter = 0;
    for (int pair = team; pair < n_pairs; pair += n_teams, ++piter) {
      const int nab = min(2, p.n_clips - 2 * pair);
      const long long qoff = static_cast<long long>(piter & 1) * p.n_slots;
      // ---- layer 0: conv0 1->C 3x3 pad 1 + ReLU in FP32 (model.py:107-111); z0 = x0 -> Q ----
      for (int ab = 0; ab < nab; ++ab) {
        const int clip = 2 * pair + ab, slot = 2 * team + ab;
        epi_bar();
        const float* src = p.mfcc + static_cast<long long>(clip) * p.H * p.W;
        for (int i = etid; i < (p.H + 2) * XS; i += T2_EPI_THREADS) {
          const int y = i / XS - 1, x = i % XS - 1;
          xs[i] = (y >= 0 && y < p.H && x >= 0 && x < p.W) ? __ldg(src + y * p.W + x) : 0.f;
        }
        epi_bar();
        const int c0 = rank * p.W / T2_G, c1 = (rank + 1) * p.W / T2_G;
        unsigned char* dst = p.Q + (qoff + slot) * p.slot_stride;
        if (valid) {
          for (int w = c0; w < c1; ++w) {
            float a[HALF];
#pragma unroll
            for (int k = 0; k < HALF; ++k) a[k] = 0.f;
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
              const float v = xs[(h + tap / 3) * XS + w + tap % 3];
#pragma unroll
              for (int k = 0; k < HALF; ++k) a[k] = fmaf(v, w0s[tap * NB + half * HALF + k], a[k]);
            }
#pragma unroll
            for (int c = 0; c < HALF / 8; ++c) {
              uint4 o;
              o.x = pack_h2(fmaxf(a[c * 8 + 0], 0.f), fmaxf(a[c * 8 + 1], 0.f));
              o.y = pack_h2(fmaxf(a[c * 8 + 2], 0.f), fmaxf(a[c * 8 + 3], 0.f));
              o.z = pack_h2(fmaxf(a[c * 8 + 4], 0.f), fmaxf(a[c * 8 + 5], 0.f));
              o.w = pack_h2(fmaxf(a[c * 8 + 6], 0.f), fmaxf(a[c * 8 + 7], 0.f));
              stcg16(dst + static_cast<long long>(w) * p.col_bytes + (half * (HALF / 8) + c) * plane_bytes + (T2_PAD + h) * 16, o);
            }
          }
        }
        __threadfence();
        fence_proxy_async_global();
        epi_bar();
        if (etid == 0) red_release_gpu_add(p.flags + clip, 1);
      }
      // ---- layers 1..n ----
      for (int L = 1; L <= p.n_layers; ++L) {
        const T2Layer* ly = p.layers + (L - 1);
        const int d = __ldg(&ly->d), flags = __ldg(&ly->flags);
        const bool add_res = flags & 1, last = flags & 2;
        const int nseg = __ldg(&ly->n_seg[rank]);
        float cst[HALF];
        {
          const float4* c4 = reinterpret_cast<const float4*>(p.cst + __ldg(&ly->cst_off) + half * HALF);
#pragma unroll
          for (int k = 0; k < HALF / 4; ++k) {
            const float4 v = __ldg(c4 + k);
            cst[4 * k] = v.x; cst[4 * k + 1] = v.y; cst[4 * k + 2] = v.z; cst[4 * k + 3] = v.w;
          }
        }
        for (int ab = 0; ab < nab; ++ab) {
          const int clip = 2 * pair + ab, slot = 2 * team + ab;
          unsigned char* Qs = p.Q + (qoff + slot) * p.slot_stride;
          unsigned char* Ps = p.P + static_cast<long long>(slot) * p.slot_stride;
          unsigned char* outb = (L & 1) ? Ps : Qs;            // odd: Q -> P ; even: P (+Q residual) -> Q in place
          float psum[HALF];
#pragma unroll
          for (int k = 0; k < HALF; ++k) psum[k] = 0.f;
          for (int s = 0; s < nseg; ++s) {
            const T2Seg* sg = &ly->seg[rank][s];
            const int r = __ldg(&sg->r), j0 = __ldg(&sg->j0), n_out = __ldg(&sg->n_out);
            for (int o = j0; o < j0 + n_out; ++o, ++g) {
              const int w = r + o * d;
              const int sl = g % NSLOT;
              const long long off = static_cast<long long>(w) * p.col_bytes + (half * (HALF / 8)) * plane_bytes + (T2_PAD + h) * 16;
              uint4 rv[HALF / 8];
              if (add_res && valid) {
#pragma unroll
                for (int c = 0; c < HALF / 8; ++c) rv[c] = ldcg16(Qs + off + c * plane_bytes);
              }
              mbar_wait(smem_u32(&bar_acc_full[sl]), (g / NSLOT) & 1);
              fence_after();
              uint32_t v[HALF];
              const uint32_t taddr = tmem + (static_cast<uint32_t>(quad * 32) << 16) + sl * NB + half * HALF;
              if constexpr (HALF == 24) {
                tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
                tmem_ld8(taddr + 16, *reinterpret_cast<uint32_t(*)[8]>(&v[16]));
              } else {
                tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
              }
              tmem_ld_wait();
              fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(smem_u32(&bar_acc_empty[sl]));
#pragma unroll
              for (int c = 0; c < HALF / 8; ++c) {
                float x[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = fmaxf(__uint_as_float(v[c * 8 + k]), 0.f) + cst[c * 8 + k];
                if (add_res && valid) {
                  const __half2* rh = reinterpret_cast<const __half2*>(&rv[c]);
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    const float2 f = __half22float2(rh[k]);
                    x[2 * k] += f.x;
                    x[2 * k + 1] += f.y;
                  }
                }
                if (!last) {
                  if (valid) {
                    uint4 ov;
                    ov.x = pack_h2(x[0], x[1]); ov.y = pack_h2(x[2], x[3]); ov.z = pack_h2(x[4], x[5]); ov.w = pack_h2(x[6], x[7]);
                    stcg16(outb + off + c * plane_bytes, ov);
                  }
                } else if (valid) {
#pragma unroll
                  for (int k = 0; k < 8; ++k) psum[c * 8 + k] += x[k];
                }
              }
            }
          }
          if (last) {
#pragma unroll
            for (int k = 0; k < HALF; ++k) {
              float sum = psum[k];
#pragma unroll
              for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
              if (lane == 0) atomicAdd(p.pool + static_cast<long long>(clip) * NB + half * HALF + k, sum);
            }
          } else {
            __threadfence();
            fence_proxy_async_global();
            epi_bar();
            if (etid == 0) red_release_gpu_add(p.flags + clip, 1);
          }
        }
      }
    }
  }

  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// logits = W'(pool / HW) + b' with the last BatchNorm folded into W', b'
__global__ void tc2_tail_kernel(const float* __restrict__ pool, const float* __restrict__ wout,
                                const float* __restrict__ bout, float* __restrict__ logits, int n_clips, int N, int L,
                                float inv_hw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_clips * L) return;
  const int clip = i / L, l = i - clip * L;
  float acc = 0.f;
  for (int k = 0; k < N; ++k) acc = fmaf(pool[clip * N + k] * inv_hw, __ldg(wout + l * N + k), acc);
  logits[i] = acc + __ldg(bout + l);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct ResTc2Plan {
  int ncg = 6, NB = 48, C = 45, H = 101, W = 40, n_layers = 13, n_labels = 12;
  int plane_rows = 117, col_bytes = 0, stage_stride = 0, wbytes = 0, smem_bytes = 0;
  long long slot_stride = 0;
  int max_teams = 37;
  T2Layer* d_layers = nullptr;
  unsigned char* d_wpack = nullptr;
  float* d_cst = nullptr;
  float* d_w0 = nullptr;
  float* d_wout = nullptr;
  float* d_bout = nullptr;
  std::vector<void*> owned;
};

void res_tc2_plan_free(ResTc2Plan* p) {
  if (!p) return;
  for (void* q : p->owned) cudaFree(q);
  delete p;
}

namespace {

template <typename T>
int up(ResTc2Plan* p, const std::vector<T>& h, T** d) {
  void* ptr = nullptr;
  HONK_CUDA(cudaMalloc(&ptr, std::max<size_t>(h.size(), 1) * sizeof(T)));
  p->owned.push_back(ptr);
  HONK_CUDA(cudaMemcpy(ptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = static_cast<T*>(ptr);
  return HONK_OK;
}

// Distributes the chains of dilation d over the T2_G CTAs of a team: whole chains by longest-
// processing-time first; when there are fewer chains than CTAs each chain is cut into equal runs.
bool plan_segments(T2Layer& ly, int W, int d) {
  for (int k = 0; k < T2_G; ++k) ly.n_seg[k] = 0;
  const int n_chain = std::min(d, W);
  auto add = [&](int rank, int r, int j0, int n_out, int len) {
    if (ly.n_seg[rank] >= T2_MAXSEG) return false;
    T2Seg& s = ly.seg[rank][ly.n_seg[rank]++];
    s.r = static_cast<short>(r);
    s.j0 = static_cast<short>(j0);
    s.n_out = static_cast<short>(n_out);
    const int a = std::max(j0 - 1, 0), b = std::min(j0 + n_out, len - 1);
    s.jin0 = static_cast<short>(a);
    s.n_in = static_cast<short>(b - a + 1);
    s.pad[0] = s.pad[1] = s.pad[2] = 0;
    return true;
  };
  if (n_chain < T2_G) {
    const int pieces = T2_G / n_chain;             // n_chain is 1 or 2 here
    for (int r = 0; r < n_chain; ++r) {
      const int len = (W - r + d - 1) / d;
      for (int k = 0; k < pieces; ++k) {
        const int a = len * k / pieces, b = len * (k + 1) / pieces;
        if (b > a && !add(r * pieces + k, r, a, b - a, len)) return false;
      }
    }
    return true;
  }
  std::vector<std::pair<int, int>> chains;          // (len, r)
  for (int r = 0; r < n_chain; ++r) chains.push_back({(W - r + d - 1) / d, r});
  std::stable_sort(chains.begin(), chains.end(), [](auto& x, auto& y) { return x.first > y.first; });
  int load[T2_G] = {0, 0, 0, 0};
  for (auto& c : chains) {
    int best = 0;
    for (int k = 1; k < T2_G; ++k)
      if (load[k] < load[best]) best = k;
    if (!add(best, c.second, 0, c.first, c.first)) return false;
    load[best] += c.first;
  }
  return true;
}

}  // namespace

int res_tc2_plan_create(honk_model* m, ResTc2Plan** out) {
  *out = nullptr;
  const honk_model_config& c = m->cfg;
  const int C = c.n_feature_maps;
  if (c.family != HONK_FAMILY_RES || C > 48 || c.res_pool[0] > 0) return HONK_OK;
  if (m->H0 > 112 || m->H0 < 8 || m->W0 < T2_G || m->W0 > 1024 || c.n_layers < 1) return HONK_OK;
  for (int i = 1; i <= c.n_layers; ++i)
    if ((c.use_dilation ? (1 << ((i - 1) / 3)) : 1) > T2_PAD) return HONK_OK;
  std::unique_ptr<ResTc2Plan, void (*)(ResTc2Plan*)> p(new ResTc2Plan(), res_tc2_plan_free);
  p->C = C;
  p->ncg = C <= 32 ? 4 : 6;
  p->NB = p->ncg * 8;
  p->H = m->H0;
  p->W = m->W0;
  p->n_layers = c.n_layers;
  p->n_labels = c.n_labels;
  p->plane_rows = p->H + T2_PAD;
  p->col_bytes = p->ncg * p->plane_rows * 16;
  p->slot_stride = align_up(static_cast<int64_t>(p->W) * p->col_bytes + T2_PAD * 16, 256);
  // a stage holds planes 0..ncg-1 of one column; the kh taps over-read up to 127 + 2*16 rows of the last plane
  p->stage_stride = static_cast<int>(align_up((p->ncg - 1) * p->plane_rows * 16 + (128 + 2 * T2_PAD) * 16, 128));
  p->wbytes = 3 * (p->ncg / 2) * 6 * p->NB * 16;
  p->smem_bytes = T2_NSTAGE * p->stage_stride + 2 * p->wbytes + ((p->H + 2) * (p->W + 3) + 9 * p->NB) * 4;
  int max_smem = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
  if (p->smem_bytes + 2048 > max_smem) return HONK_OK;
  p->max_teams = std::max(1, m->n_sms / T2_G);
  const int N = p->NB, ncg = p->ncg, NKC = ncg / 2;

  // per-layer BatchNorm statistics (mean_0 = 0, scale_0 = 1: conv1 reads the raw conv0 output)
  std::vector<std::vector<float>> mean(c.n_layers + 1, std::vector<float>(N, 0.f)), scale(c.n_layers + 1, std::vector<float>(N, 1.f));
  for (int i = 1; i <= c.n_layers; ++i) {
    const HostTensor& mu = m->host.at("bn" + std::to_string(i) + ".running_mean");
    const HostTensor& var = m->host.at("bn" + std::to_string(i) + ".running_var");
    for (int k = 0; k < C; ++k) {
      mean[i][k] = mu.data[k];
      scale[i][k] = 1.0f / std::sqrt(var.data[k] + 1e-5f);       // BatchNorm2d(affine=False), eval
    }
  }
  std::vector<T2Layer> layers(c.n_layers);
  std::vector<unsigned char> wpack(static_cast<size_t>(c.n_layers) * p->wbytes, 0);
  std::vector<float> cst(static_cast<size_t>(c.n_layers) * N, 0.f);
  for (int L = 1; L <= c.n_layers; ++L) {
    T2Layer& ly = layers[L - 1];
    memset(&ly, 0, sizeof(ly));
    ly.d = c.use_dilation ? (1 << ((L - 1) / 3)) : 1;
    const bool even = (L % 2 == 0), last = (L == c.n_layers);
    ly.flags = (even ? 1 : 0) | (last ? 2 : 0);
    ly.w_off = (L - 1) * p->wbytes;
    ly.cst_off = (L - 1) * N;
    if (!plan_segments(ly, p->W, ly.d)) return HONK_OK;
    // epilogue constant: z_L = relu(acc) [+ z_{L-2}] + cst ; last layer pools x_L itself
    for (int k = 0; k < C; ++k) {
      float v = even ? mean[L - 2][k] : 0.f;
      if (!last) v -= mean[L][k];
      cst[static_cast<size_t>(L - 1) * N + k] = v;
    }
    // B operand: [i = kh*NKC + kc][K half][n = blk*NB + co][ci % 8], blk <-> kw = 2 - blk,
    // value = W[co][ci][kh][kw] * scale_{L-1}[ci]
    const HostTensor& w = m->host.at("conv" + std::to_string(L) + ".weight");
    __half* hw = reinterpret_cast<__half*>(wpack.data() + static_cast<size_t>(L - 1) * p->wbytes);
    for (int kh = 0; kh < 3; ++kh)
      for (int kc = 0; kc < NKC; ++kc)
        for (int khalf = 0; khalf < 2; ++khalf)
          for (int blk = 0; blk < 3; ++blk)
            for (int co = 0; co < C; ++co)
              for (int e = 0; e < 8; ++e) {
                const int ci = kc * 16 + khalf * 8 + e;
                if (ci >= C) continue;
                const int kw = 2 - blk;
                const float v = w.data[((static_cast<size_t>(co) * C + ci) * 3 + kh) * 3 + kw] * scale[L - 1][ci];
                const size_t idx = ((static_cast<size_t>(kh * NKC + kc) * 2 + khalf) * (3 * N) + blk * N + co) * 8 + e;
                hw[idx] = __float2half(v);
              }
  }
  HONK_TRY(up(p.get(), layers, &p->d_layers));
  HONK_TRY(up(p.get(), wpack, &p->d_wpack));
  HONK_TRY(up(p.get(), cst, &p->d_cst));
  {
    const HostTensor& w = m->host.at("conv0.weight");
    std::vector<float> h(9 * N, 0.f);
    for (int co = 0; co < C; ++co)
      for (int t = 0; t < 9; ++t) h[t * N + co] = w.data[co * 9 + t];
    HONK_TRY(up(p.get(), h, &p->d_w0));
  }
  {
    // logits = W_o (s (.) (mean_hw(x_n) - mu_n)) + b_o  (last BN commutes with the mean)
    const HostTensor& w = m->host.at("output.weight");
    const HostTensor& b = m->host.at("output.bias");
    std::vector<float> hwv(static_cast<size_t>(c.n_labels) * N, 0.f), hb(c.n_labels, 0.f);
    for (int l = 0; l < c.n_labels; ++l) {
      double acc = b.data[l];
      for (int k = 0; k < C; ++k) {
        const float ws = w.data[static_cast<size_t>(l) * C + k] * scale[c.n_layers][k];
        hwv[static_cast<size_t>(l) * N + k] = ws;
        acc -= static_cast<double>(ws) * mean[c.n_layers][k];
      }
      hb[l] = static_cast<float>(acc);
    }
    HONK_TRY(up(p.get(), hwv, &p->d_wout));
    HONK_TRY(up(p.get(), hb, &p->d_bout));
  }
  if (ncg == 6) HONK_CUDA(cudaFuncSetAttribute(tc2_res_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes));
  else HONK_CUDA(cudaFuncSetAttribute(tc2_res_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes));
  *out = p.release();
  return HONK_OK;
}

static int tc2_teams(const ResTc2Plan* p, int64_t n_clips) {
  return static_cast<int>(std::min<int64_t>(p->max_teams, (n_clips + 1) / 2));
}

int64_t res_tc2_workspace_bytes(const ResTc2Plan* p, int64_t n_clips) {
  const int n_slots = 2 * tc2_teams(p, n_clips);
  Arena a(nullptr, 0);
  a.take<unsigned char>(3ll * n_slots * p->slot_stride);
  a.take<float>(n_clips * p->NB);
  a.take<unsigned int>(n_clips);
  return align_up(a.off, 256);
}

int res_tc2_forward(const ResTc2Plan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                    cudaStream_t s) {
  const int teams = tc2_teams(p, n_clips), n_slots = 2 * teams;
  Arena ar(ws, ws_bytes);
  unsigned char* act = ar.take<unsigned char>(3ll * n_slots * p->slot_stride);
  float* pool = ar.take<float>(n_clips * p->NB);
  unsigned int* flags = ar.take<unsigned int>(n_clips);
  HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: tcgen05 workspace overrun");
  // zero rows between planes must be zero; pool sums and layer counters start at zero
  HONK_CUDA(cudaMemsetAsync(act, 0, static_cast<size_t>(ar.off), s));
  T2Params a{};
  a.layers = p->d_layers;
  a.wpack = p->d_wpack;
  a.cst = p->d_cst;
  a.mfcc = d_mfcc;
  a.w0 = p->d_w0;
  a.P = act;
  a.Q = act + static_cast<long long>(n_slots) * p->slot_stride;
  a.pool = pool;
  a.flags = flags;
  a.slot_stride = p->slot_stride;
  a.n_layers = p->n_layers;
  a.n_clips = static_cast<int>(n_clips);
  a.n_slots = n_slots;
  a.H = p->H;
  a.W = p->W;
  a.plane_rows = p->plane_rows;
  a.col_bytes = p->col_bytes;
  a.stage_stride = p->stage_stride;
  a.wbytes = p->wbytes;
  {
    ProfSpan span("res_conv", s);
    if (p->ncg == 6) tc2_res_kernel<6><<<teams * T2_G, T2_THREADS, p->smem_bytes, s>>>(a);
    else tc2_res_kernel<4><<<teams * T2_G, T2_THREADS, p->smem_bytes, s>>>(a);
    HONK_LAUNCH_CHECK();
  }
  {
    ProfSpan span("res_tail", s);
    const int total = static_cast<int>(n_clips) * p->n_labels;
    tc2_tail_kernel<<<(total + 127) / 128, 128, 0, s>>>(pool, p->d_wout, p->d_bout, d_logits, static_cast<int>(n_clips), p->NB,
                                                        p->n_labels, 1.0f / static_cast<float>(p->H * p->W));
    HONK_LAUNCH_CHECK();
  }
  return HONK_OK;
}

}  // namespace honk
