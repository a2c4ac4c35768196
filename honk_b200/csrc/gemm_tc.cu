// Generic tcgen05 path for the Google cnn-* family (SpeechModel.forward, utils/model.py:186-205) other than the
// cnn-trad-pool2 geometry (which has its own fused kernel, cnn_tc.cu): every layer - conv1, conv2 (strided, pooled,
// full-height kernels), lin, dnn1, dnn2, output (model.py:385-407 configs) - is ONE GEMM on the tensor cores,
//     C[M][N] = act(A[M][K] B[N][K]^T + bias),     f16 operands, FP32 accumulators in TMEM,
// fed by a pack kernel that writes the layer's A operand (im2col rows, with the previous layer's MaxPool folded in) in the
// tiled layout the MMA reads:
//     A: [M tile of 128 rows][K plane of 8 elements][128 rows][8 halves]   (one 16-byte unit per row and plane; a stage of
//        KC planes of one tile is one contiguous bulk copy; UMMA K-major no-swizzle: SBO = 128 B, LBO = one plane)
//     B: [N tile][K plane][NT rows][8 halves], packed once at model load.
// Row m of A = (clip, output position), k = (kh * KW + kw) * Cs + c over the NHWC f16 activation of the previous layer
// (conv1: k = kh * KW + kw over the float32 MFCC image).  Linear layers are the 1 x 1 case over a 1 x 1 map with C = K; their
// weights are permuted at load from the reference's (C, H, W) flatten order (model.py:195) to the (H, W, C) order stored here.
//
// gemm_tc_kernel: persistent, one CTA per SM over (M tile, N tile) pairs.  warp 0 = bulk-copy producer (4 stages of KC = 8
// planes: 16 KB of A + NT * 128 B of B), warp 1 = MMA issuer (M = 128, N = NT <= 256, K = 16 per tcgen05.mma, two TMEM
// accumulator buffers), warps 2-5 = epilogue (TMEM -> + bias -> ReLU -> f16 / f32 rows in global memory).
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

constexpr int GT_THREADS = 192;
constexpr int GT_KC = 8;                 // K planes (of 8 elements) per stage
constexpr int GT_NSTAGE = 4;
constexpr int GT_MAXNT = 256;

struct GemmTcParams {
  const unsigned char* A;      // tiled, see above
  const unsigned char* B;
  const float* bias;           // [n_ntiles * NT] (zero in padded columns)
  void* C;                     // row-major [M][ldc], f16 or f32
  long long M;
  int KP;                      // K planes (multiple of 2)
  int NT, n_ntiles;            // N tile (multiple of 16, <= 256) and count
  int ldc;                     // elements per output row
  int n_store;                 // columns of C that exist (<= ldc); columns >= n_store are not written
  int relu, out_f32;
  long long n_mtiles;
  // implicit mode (stride-1 convolutions over a planar activation, no materialised A): A plane (tap, cg) of the tile of clip
  // `clip` starting at flat input row r0 is the 128 rows [r0 + kh * Wp + kw, +128) of plane cg of that clip's image
  //   Ap[clip][cg][R_alloc rows][8 halves],   row = h * Wp + w over the (pooled) input map.
  // Output row r0 + i is output position (ho, wo) = ((r0 + i) / Wp, (r0 + i) % Wp); rows with wo >= Wo or ho >= Ho are junk.
  const unsigned char* Ap;
  int imp, NCG, R_alloc, Wp, KW, n_taps, tiles_per_clip, Ho, Wo;
};

__device__ __forceinline__ uint32_t gt_elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
__device__ __forceinline__ void gt_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ uint32_t gt_pack_h2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

__global__ void __launch_bounds__(GT_THREADS, 1) gemm_tc_kernel(const __grid_constant__ GemmTcParams p) {
  extern __shared__ __align__(1024) unsigned char gt_smem[];
  __shared__ __align__(8) uint64_t bar_full[GT_NSTAGE], bar_empty[GT_NSTAGE], bar_accfull[2], bar_accempty[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int a_stage = GT_KC * 2048, b_stage = GT_KC * p.NT * 16, stage_bytes = a_stage + b_stage;
  const uint32_t a_full = smem_u32(&bar_full[0]), a_empty = smem_u32(&bar_empty[0]);
  const uint32_t a_accfull = smem_u32(&bar_accfull[0]), a_accempty = smem_u32(&bar_accempty[0]);
  if (tid == 0) {
    for (int s = 0; s < GT_NSTAGE; ++s) {
      mbar_init(a_full + 8 * s, 1);
      mbar_init(a_empty + 8 * s, 1);
    }
    for (int k = 0; k < 2; ++k) {
      mbar_init(a_accfull + 8 * k, 1);
      mbar_init(a_accempty + 8 * k, 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const long long n_tiles = p.n_mtiles * p.n_ntiles;
  const int n_ks = (p.KP + GT_KC - 1) / GT_KC;                      // stages per tile

  if (warp == 0) {
    // ======================================= producer =======================================
    // lane 0 arms the stage's barrier and copies B (and A in the materialised mode); in implicit mode lane i copies plane i of
    // the stage (8 bulk copies of 2 KB issued side by side instead of one lane walking them)
    uint32_t it = 0;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      const long long mt = t / p.n_ntiles;
      const int nt = static_cast<int>(t - mt * p.n_ntiles);
      const unsigned char* At = p.A + mt * (static_cast<long long>(p.KP) * 2048);
      const unsigned char* Bt = p.B + static_cast<long long>(nt) * p.KP * p.NT * 16;
      const long long clip = p.imp ? mt / p.tiles_per_clip : 0;
      const unsigned char* Ac = p.imp ? p.Ap + (clip * p.NCG * p.R_alloc + (mt - clip * p.tiles_per_clip) * 128) * 16 : nullptr;
      for (int ks = 0; ks < n_ks; ++ks, ++it) {
        const uint32_t s = it % GT_NSTAGE;
        const int kc = min(GT_KC, p.KP - ks * GT_KC);
        if (lane == 0) {
          if (it >= GT_NSTAGE) mbar_wait(a_empty + 8 * s, ((it / GT_NSTAGE) & 1) ^ 1);
          mbar_expect_tx(a_full + 8 * s, kc * 2048 + kc * p.NT * 16);
          if (!p.imp) gt_bulk_load(smem_u32(gt_smem + s * stage_bytes), At + static_cast<long long>(ks) * GT_KC * 2048, kc * 2048, a_full + 8 * s);
          gt_bulk_load(smem_u32(gt_smem + s * stage_bytes + a_stage), Bt + static_cast<long long>(ks) * GT_KC * p.NT * 16, kc * p.NT * 16,
                       a_full + 8 * s);
        }
        if (p.imp) {
          __syncwarp();                                            // the barrier is armed and the stage is free
          if (lane < kc) {
            const int kp = ks * GT_KC + lane, tap = kp / p.NCG, cg = kp - tap * p.NCG, kh = tap / p.KW, kw = tap - kh * p.KW;
            const unsigned char* src = tap < p.n_taps ? Ac + (static_cast<long long>(cg) * p.R_alloc + kh * p.Wp + kw) * 16 : Ac;   // padding plane: zero weights
            gt_bulk_load(smem_u32(gt_smem + s * stage_bytes + lane * 2048), src, 2048, a_full + 8 * s);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer =====================================
    const uint32_t idesc = make_idesc(128, p.NT, 0);
    const uint64_t a_hi = make_desc(0, 2048, 128) & 0xFFFFFFFFFFFF0000ull;             // LBO = one plane of 128 rows
    const uint64_t b_hi = make_desc(0, p.NT * 16, 128) & 0xFFFFFFFFFFFF0000ull;
    uint32_t it = 0, c = 0;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++c) {
      const uint32_t ab = c & 1;
      if (c >= 2) {
        mbar_wait(a_accempty + 8 * ab, ((c >> 1) & 1) ^ 1);
        fence_after();
      }
      const uint32_t d = tmem + ab * GT_MAXNT;
      for (int ks = 0; ks < n_ks; ++ks, ++it) {
        const uint32_t s = it % GT_NSTAGE;
        mbar_wait(a_full + 8 * s, (it / GT_NSTAGE) & 1);
        fence_after();
        const int kc = min(GT_KC, p.KP - ks * GT_KC);
        if (gt_elect_one()) {
          const uint32_t sa = smem_u32(gt_smem + s * stage_bytes), sb = sa + a_stage;
          for (int j = 0; j < kc / 2; ++j) {
            const uint64_t ad = a_hi | ((sa + j * 2 * 2048) >> 4);
            const uint64_t bd = b_hi | ((sb + j * 2 * p.NT * 16) >> 4);
            mma_f16_ss(d, ad, bd, idesc, (ks > 0 || j > 0) ? 1u : 0u);
          }
          commit(a_empty + 8 * s);
          if (ks == n_ks - 1) commit(a_accfull + 8 * ab);
        }
        __syncwarp();
      }
    }
  } else {
    // ======================================= epilogue =======================================
    const int quad = warp & 3, row = quad * 32 + lane;
    const uint32_t tlane = static_cast<uint32_t>(quad * 32) << 16;
    uint32_t c = 0;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++c) {
      const long long mt = t / p.n_ntiles;
      const int nt = static_cast<int>(t - mt * p.n_ntiles);
      const uint32_t ab = c & 1;
      mbar_wait(a_accfull + 8 * ab, (c >> 1) & 1);
      fence_after();
      long long m = mt * 128 + row;
      if (p.imp) {                                                 // flat input row -> output position; junk rows are not stored
        const long long clip = mt / p.tiles_per_clip;
        const int r = static_cast<int>(mt - clip * p.tiles_per_clip) * 128 + row, ho = r / p.Wp, wo = r - ho * p.Wp;
        m = (ho < p.Ho && wo < p.Wo) ? (clip * p.Ho + ho) * p.Wo + wo : p.M;
      }
      const int col0 = nt * p.NT;
      for (int c0 = 0; c0 < p.NT; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + tlane + ab * GT_MAXNT + c0, v);
        tmem_ld_wait();
        if (c0 + 16 >= p.NT) {                                     // the accumulator is in registers: hand it back
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(a_accempty + 8 * ab);
        }
        if (m >= p.M || col0 + c0 >= p.n_store) continue;
        float f[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          f[k] = __uint_as_float(v[k]) + __ldg(p.bias + col0 + c0 + k);
          if (p.relu) f[k] = fmaxf(f[k], 0.f);
        }
        if (p.out_f32) {
          float* dst = static_cast<float*>(p.C) + m * p.ldc + col0 + c0;
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (col0 + c0 + k < p.n_store) dst[k] = f[k];
        } else {
          __half* dst = static_cast<__half*>(p.C) + m * p.ldc + col0 + c0;
          if (col0 + c0 + 16 <= p.n_store && (p.ldc & 7) == 0) {
            *reinterpret_cast<uint4*>(dst) = make_uint4(gt_pack_h2(f[0], f[1]), gt_pack_h2(f[2], f[3]), gt_pack_h2(f[4], f[5]), gt_pack_h2(f[6], f[7]));
            *reinterpret_cast<uint4*>(dst + 8) =
                make_uint4(gt_pack_h2(f[8], f[9]), gt_pack_h2(f[10], f[11]), gt_pack_h2(f[12], f[13]), gt_pack_h2(f[14], f[15]));
          } else {
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (col0 + c0 + k < p.n_store) dst[k] = __float2half_rn(fminf(fmaxf(f[k], -65504.f), 65504.f));
          }
        }
      }
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}


// Implicit-mode convolution with the clip's planar image RESIDENT in shared memory (the stride-1 second convolutions of
// the two-convolution cnn-* configs).  A persistent CTA takes whole clips: the producer copies the clip's NCG planes
// (R_alloc rows x 16 B each) once, then streams the layer's B tiles through a ring; the MMA warp runs, for every B stage, the
// stage's K = 16 steps for EVERY M tile of the clip (all tiles_per_clip accumulators live in TMEM), reading the A operand as
// tap-shifted windows of the resident image: rows [t * 128 + kh * Wp + kw, +128) of planes (cg, cg + 1).  Per clip the L2 is
// asked for the image once and the weights once instead of (taps x tiles) shifted copies of the image.
constexpr int CR_EPI = 16;               // epilogue warps: four per TMEM lane quarter, a quarter of the channels each
constexpr int CR_THREADS = (2 + CR_EPI) * 32;
constexpr int CR_MAXSTEPS = 1024;     // K = 16 steps of a layer (taps x channel-group pairs)
constexpr int CR_MAXSTAGE = 12;       // B stages in flight: as many as the shared memory left by the image holds (bulk-copy latency)

struct ConvResParams {
  const unsigned char* Ap;     // [n_clips][NCG][R_alloc][16 B]
  const unsigned char* B;      // [KP][NT][8 halves], KP = n_taps * NCG
  const float* bias;
  __half* C;                   // NHWC f16 [n_clips][Ho][Wo][ldc]
  int n_clips, NCG, R_alloc, Wp, KW, n_taps, tiles_per_clip, Ho, Wo, NT, ldc, relu, n_stage;
  int mode;                    // HONK_CR_MODE, timing experiments only (results are garbage): 1 no image reloads, 2 no epilogue math / stores, 4 no MMAs
  int G;                       // clips per pass: their images are resident together and share one pass of the weight stream
};

template <int TPC>       // M tiles per clip as a constant (0: p.tiles_per_clip), so the issuing warp's inner loop unrolls
__global__ void __launch_bounds__(CR_THREADS, 1) conv_res_kernel(const __grid_constant__ ConvResParams p) {
  extern __shared__ __align__(1024) unsigned char cr_smem[];
  __shared__ __align__(8) uint64_t bar_bfull[CR_MAXSTAGE], bar_bempty[CR_MAXSTAGE], bar_img[2], bar_acc[2];
  __shared__ uint32_t a_off[CR_MAXSTEPS];                         // byte offset of every K = 16 step's A window inside the image
  const uint32_t CR_NSTAGE = p.n_stage;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int plane_bytes = p.R_alloc * 16, img_bytes = p.NCG * plane_bytes, b_stage = GT_KC * p.NT * 16;
  unsigned char* ring = cr_smem + (p.G * img_bytes + 1023) / 1024 * 1024;
  const int n_pass = (p.n_clips + p.G - 1) / p.G, tpc = TPC ? TPC : p.tiles_per_clip;
  const uint32_t a_bfull = smem_u32(&bar_bfull[0]), a_bempty = smem_u32(&bar_bempty[0]);
  const uint32_t a_imgfull = smem_u32(&bar_img[0]), a_imgempty = smem_u32(&bar_img[1]);
  const uint32_t a_accfull = smem_u32(&bar_acc[0]), a_accempty = smem_u32(&bar_acc[1]);
  if (tid == 0) {
    for (int s = 0; s < CR_MAXSTAGE; ++s) {
      mbar_init(a_bfull + 8 * s, 1);
      mbar_init(a_bempty + 8 * s, 1);
    }
    mbar_init(a_imgfull, 1);
    mbar_init(a_imgempty, 1);
    mbar_init(a_accfull, 1);
    mbar_init(a_accempty, CR_EPI);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  const int KP = p.n_taps * p.NCG, n_ks = (KP + GT_KC - 1) / GT_KC;
  for (int st = tid; st < KP / 2; st += CR_THREADS) {               // the issuing warp is bound by its own instruction stream:
    const int tap = (2 * st) / p.NCG, cg = 2 * st - tap * p.NCG, kh = tap / p.KW, kw = tap - kh * p.KW;   // no index arithmetic there
    a_off[st] = static_cast<uint32_t>(cg * (p.R_alloc * 16) + (kh * p.Wp + kw) * 16);
  }
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;

  if (warp == 0) {
    // ======================================= producer =======================================
    if (lane == 0) {
      uint32_t it = 0, n = 0;
      for (int q = blockIdx.x; q < n_pass; q += gridDim.x, ++n) {
        if (n > 0) mbar_wait(a_imgempty, (n - 1) & 1);             // the previous pass's MMAs have read the images
        const int n_here = min(p.G, p.n_clips - q * p.G);
        const bool skip_img = (p.mode & 1) && n > 0;
        mbar_expect_tx(a_imgfull, skip_img ? plane_bytes : n_here * img_bytes);
        const unsigned char* src = p.Ap + static_cast<long long>(q) * p.G * img_bytes;   // the pass's clips are contiguous
        for (int pl = 0; pl < (skip_img ? 1 : n_here * p.NCG); ++pl)
          gt_bulk_load(smem_u32(cr_smem + pl * plane_bytes), src + static_cast<long long>(pl) * plane_bytes, plane_bytes, a_imgfull);
        for (int ks = 0; ks < n_ks; ++ks, ++it) {
          const uint32_t s = it % CR_NSTAGE;
          if (it >= CR_NSTAGE) mbar_wait(a_bempty + 8 * s, ((it / CR_NSTAGE) & 1) ^ 1);
          const int kc = min(GT_KC, KP - ks * GT_KC);
          mbar_expect_tx(a_bfull + 8 * s, kc * p.NT * 16);
          gt_bulk_load(smem_u32(ring + s * b_stage), p.B + static_cast<long long>(ks) * GT_KC * p.NT * 16, kc * p.NT * 16, a_bfull + 8 * s);
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer =====================================
    const uint32_t idesc = make_idesc(128, p.NT, 0);
    const uint64_t a_hi = make_desc(0, plane_bytes, 128) & 0xFFFFFFFFFFFF0000ull;      // LBO = one channel-group plane
    const uint64_t b_hi = make_desc(0, p.NT * 16, 128) & 0xFFFFFFFFFFFF0000ull;
    const uint32_t img = smem_u32(cr_smem);
    uint32_t it = 0, n = 0;
    for (int q = blockIdx.x; q < n_pass; q += gridDim.x, ++n) {
      const int n_g = (p.mode & 4) ? 0 : min(p.G, p.n_clips - q * p.G);
      mbar_wait(a_imgfull, n & 1);
      if (n > 0) mbar_wait(a_accempty, (n - 1) & 1);               // the epilogue has drained the previous pass's accumulators
      fence_after();
      for (int ks = 0; ks < n_ks; ++ks, ++it) {
        const uint32_t s = it % CR_NSTAGE;
        mbar_wait(a_bfull + 8 * s, (it / CR_NSTAGE) & 1);
        fence_after();
        const int kc = min(GT_KC, KP - ks * GT_KC);
        if (gt_elect_one()) {
          const uint32_t sb = smem_u32(ring + s * b_stage);
          for (int j = 0; j < kc / 2; ++j) {
            uint32_t a_g = (img + a_off[ks * (GT_KC / 2) + j]) >> 4, d_g = tmem;
            const uint64_t bd = b_hi | ((sb + j * 2 * p.NT * 16) >> 4);
            const uint32_t acc = (ks > 0 || j > 0) ? 1u : 0u;
            for (int g = 0; g < n_g; ++g, a_g += img_bytes >> 4, d_g += tpc * p.NT) {
#pragma unroll
              for (int t = 0; t < tpc; ++t) mma_f16_ss(d_g + t * p.NT, a_hi | (a_g + t * 128), bd, idesc, acc);
            }
          }
          commit(a_bempty + 8 * s);
          if (ks == n_ks - 1) {
            commit(a_accfull);
            commit(a_imgempty);
          }
        }
        __syncwarp();
      }
    }
  } else {
    // ======================================= epilogue =======================================
    const int quad = warp & 3, row = quad * 32 + lane, part = (warp - 2) >> 2;          // TMEM lane quarter, column part
    const uint32_t tlane = static_cast<uint32_t>(quad * 32) << 16;
    const int n16 = p.NT / 16, g_lo = part * n16 / (CR_EPI / 4), g_hi = (part + 1) * n16 / (CR_EPI / 4);   // 16-column groups of this warp
    uint32_t n = 0;
    for (int q = blockIdx.x; q < n_pass; q += gridDim.x, ++n) {
      const int n_here = min(p.G, p.n_clips - q * p.G), n_t = n_here * tpc;
      mbar_wait(a_accfull, n & 1);
      fence_after();
      if (g_lo == g_hi && lane == 0) mbar_arrive(a_accempty);        // fewer column groups than parts
      for (int tt = 0, g = 0, t = 0; tt < n_t; ++tt, g += (t == tpc - 1), t = (t == tpc - 1) ? 0 : t + 1) {
        const long long clip = static_cast<long long>(q) * p.G + g;
        const int r = t * 128 + row, ho = r / p.Wp, wo = r - ho * p.Wp;
        const bool valid = ho < p.Ho && wo < p.Wo;
        __half* dst = p.C + ((clip * p.Ho + ho) * p.Wo + wo) * p.ldc;
        for (int c0 = g_lo * 16; c0 < g_hi * 16; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tmem + tlane + tt * p.NT + c0, v);
          tmem_ld_wait();
          if (tt == n_t - 1 && c0 + 16 >= g_hi * 16) {                  // everything is in registers: hand the accumulators back
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_accempty);
          }
          if (!valid || c0 >= p.ldc || (p.mode & 2)) continue;
          float f[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            f[k] = __uint_as_float(v[k]) + __ldg(p.bias + c0 + k);
            if (p.relu) f[k] = fmaxf(f[k], 0.f);
          }
          *reinterpret_cast<uint4*>(dst + c0) = make_uint4(gt_pack_h2(f[0], f[1]), gt_pack_h2(f[2], f[3]), gt_pack_h2(f[4], f[5]), gt_pack_h2(f[6], f[7]));
          *reinterpret_cast<uint4*>(dst + c0 + 8) =
              make_uint4(gt_pack_h2(f[8], f[9]), gt_pack_h2(f[10], f[11]), gt_pack_h2(f[12], f[13]), gt_pack_h2(f[14], f[15]));
        }
      }
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------
// conv1 (1 -> C1, KH x 8 kernel, stride (sh, 1)) without an im2col operand.  A pre-pass stores each clip's MFCC image as
//   x8p[clip][phase p < sh][wx < W - 7][i < Hph][8 halves] = f16(x[i * sh + p][wx .. wx + 7])      (Hph = ceil(H / sh)),
// i.e. the 8 coefficient taps of one kernel row are one 16-byte unit and, per time phase, time runs fastest.  With output rows
// numbered r = wo * Hph + ho, kernel row kh of EVERY output row of a tile is the run of 128 units starting at
// r0 + (kh % sh) * Wx * Hph + kh / sh: the A operand is a window of the image, two kernel rows per K = 16 MMA (their runs are a
// constant distance apart: 16 bytes for sh = 1, one phase image otherwise).  A persistent CTA keeps the layer's weights and two
// clips' images resident, walks the clip's M tiles with two TMEM accumulators, and its epilogue stores + bias, ReLU, f16 NHWC.
// ------------------------------------------------------------------------------------------
struct Conv1ResParams {
  const unsigned char* X;      // x8p images, img_bytes per clip (zero padded past the last real row)
  const unsigned char* B;      // [KP planes][NT][8 halves], plane = kernel row (KP even)
  const float* bias;
  __half* C;                   // NHWC f16 [n_clips][Ho][Wo][ldc]
  int n_clips, img_bytes, KP, NT, ldc, Hph, Ho, Wo, n_tiles, lbo, sh, phase_stride;   // phase_stride = Wx * Hph rows
  int mode;                    // HONK_C1_MODE, timing experiments only (2, 4: results are garbage): 2 no epilogue stores, 4 no MMAs, 8 two accumulators
  // fused MaxPool(ph, pw) (model.py:190): Hph = 128, so tile t is coefficient column wo = t and a thread keeps ONE time row in
  // every tile; the epilogue writes the pooled map as the next layer's planar image P[clip][cg][R_alloc rows][8 halves]
  // fuse = 2 (time stride > 1): column stride Hc = 1 << hc_shift <= 64 rows, several columns per tile; the ReLU-ed f16 rows go to a
  // ring of n_cols columns in shared memory and every complete block of pw columns is pooled from there
  int fuse, ph, pw, Hp, Wp, R_alloc, cpt, hc_shift, n_cols;
  unsigned char* P;
};

__global__ void __launch_bounds__(256) x8p_pack_kernel(const float* __restrict__ mfcc, unsigned char* __restrict__ X, int H, int W, int sh, int Hph,
                                                       int img_units) {
  const int Wx = W - 7;
  const float* src = mfcc + static_cast<long long>(blockIdx.x) * H * W;
  uint4* dst = reinterpret_cast<uint4*>(X) + static_cast<long long>(blockIdx.x) * img_units;
  for (int u = threadIdx.x; u < img_units; u += blockDim.x) {
    uint4 o = make_uint4(0, 0, 0, 0);
    const int i = u % Hph, t = u / Hph, wx = t % Wx, ph = t / Wx;
    const int h = i * sh + ph;
    if (ph < sh && h < H) {
      uint32_t hv[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) hv[e] = __half_as_ushort(__float2half_rn(fminf(fmaxf(__ldg(src + h * W + wx + e), -60000.f), 60000.f)));
      o = make_uint4(hv[0] | (hv[1] << 16), hv[2] | (hv[3] << 16), hv[4] | (hv[5] << 16), hv[6] | (hv[7] << 16));
    }
    dst[u] = o;
  }
}

__device__ __forceinline__ uint32_t pk_hmax2(uint32_t a, uint32_t b) {
  uint32_t o;
  asm("max.f16x2 %0, %1, %2;" : "=r"(o) : "r"(a), "r"(b));
  return o;
}

// The same image for the 40-coefficient geometry, one thread per TIME ROW: the row's 40 values are read once (contiguous), kept as 20
// packed words, and the 33 units of the row (sliding windows of 8 halves) go out with consecutive threads = consecutive time rows of
// one coefficient column, i.e. coalesced 16-byte stores.  Rows are numbered phase-major like the image (row = phase * Hph + i).
__global__ void __launch_bounds__(128) x8p_pack_rows40_kernel(const float* __restrict__ mfcc, unsigned char* __restrict__ X, int H, int sh, int Hph,
                                                              int img_units) {
  constexpr int W = 40, Wx = 33;
  const float* src = mfcc + static_cast<long long>(blockIdx.x) * H * W;
  uint4* dst = reinterpret_cast<uint4*>(X) + static_cast<long long>(blockIdx.x) * img_units;
  for (int row = threadIdx.x; row < sh * Hph; row += blockDim.x) {
    const int ph = row / Hph, i = row - ph * Hph, h = i * sh + ph;
    uint32_t w[W / 2 + 1];
    if (h < H) {
      const float4* s4 = reinterpret_cast<const float4*>(src + h * W);                 // clips and rows are multiples of 16 bytes
#pragma unroll
      for (int k = 0; k < W / 4; ++k) {
        const float4 q = __ldg(s4 + k);
        w[2 * k] = gt_pack_h2(fminf(fmaxf(q.x, -60000.f), 60000.f), fminf(fmaxf(q.y, -60000.f), 60000.f));
        w[2 * k + 1] = gt_pack_h2(fminf(fmaxf(q.z, -60000.f), 60000.f), fminf(fmaxf(q.w, -60000.f), 60000.f));
      }
    } else {
#pragma unroll
      for (int k = 0; k < W / 2; ++k) w[k] = 0;
    }
    w[W / 2] = 0;
    uint4* d = dst + static_cast<long long>(ph) * Wx * Hph + i;
#pragma unroll
    for (int wx = 0; wx < Wx; ++wx) {
      const int k = wx >> 1;
      uint4 o;
      if (wx & 1) {
        o = make_uint4(__funnelshift_r(w[k], w[k + 1], 16), __funnelshift_r(w[k + 1], w[k + 2], 16), __funnelshift_r(w[k + 2], w[k + 3], 16),
                       __funnelshift_r(w[k + 3], w[k + 4], 16));
      } else {
        o = make_uint4(w[k], w[k + 1], w[k + 2], w[k + 3]);
      }
      d[static_cast<long long>(wx) * Hph] = o;
    }
  }
  for (int u = sh * Wx * Hph + threadIdx.x; u < img_units; u += blockDim.x) dst[u] = make_uint4(0, 0, 0, 0);   // the tiles' over-read
}

constexpr int C1_FUSE_MAXNT = 128;       // fused pool: the column maxima of a warp's half of the channels stay in registers
constexpr int C1_UNROLL = 12;            // kernel-row pairs issued from registers (KH <= 24: every reference config)
constexpr int C1_MAXSTEPS = 64;          // kernel-row pairs of conv1 (KH <= 128)
constexpr int C1_EPI = 16;               // epilogue warps: four per TMEM lane quarter, a quarter of the channels each (their dependent
                                         // instruction chains, not issue slots, bound the kernel: 8 warps measured 25-35 % slower)
constexpr int C1_ISSUERS = 2;            // MMA-issuing warps, one per TMEM accumulator (alternate tiles): a lone issuer needs ~97 clk per
                                         // MMA of its own instruction stream (ncu, profiles/r05_conv1_res_ncu.txt) against 56 clk of tensor time
constexpr int C1_EPI0 = 1 + C1_ISSUERS;  // first epilogue warp
constexpr int C1_THREADS = (C1_EPI0 + C1_EPI) * 32;      // producer, MMA issuers, epilogue

__global__ void __launch_bounds__(C1_THREADS, 1) conv1_res_kernel(const __grid_constant__ Conv1ResParams p) {
  extern __shared__ __align__(1024) unsigned char c1_smem[];
  __shared__ __align__(8) uint64_t bar_w, bar_imgfull[2], bar_imgempty[2], bar_accfull[4], bar_accempty[4];
  __shared__ __align__(16) float bias_s[GT_MAXNT];
  __shared__ uint32_t c1_aoff[C1_MAXSTEPS];                       // A window of every K = 16 step (a kernel-row pair), bytes from the tile's first unit
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int img_al = (p.img_bytes + 1023) / 1024 * 1024, w_bytes = p.KP * p.NT * 16;
  // TMEM accumulators in flight: four of 128 columns when the layer fits them, else two of 256.  The commit -> epilogue load ->
  // hand-back round trip is ~1700 clk (measured with MMAs and stores switched off), so two accumulators cap a tile at ~860 clk
  const uint32_t acc_shift = (p.NT <= 128 && !(p.mode & 8)) ? 2 : 1, acc_mask = (1u << acc_shift) - 1, acc_stride = 512u >> acc_shift;
  for (int j = tid; j < p.KP / 2; j += C1_THREADS) {               // no index arithmetic in the issuing warp's loop
    const int kh = 2 * j;                                          // kernel rows (kh, kh + 1): phases kh % sh and the next one
    c1_aoff[j] = static_cast<uint32_t>(((kh % p.sh) * p.phase_stride + kh / p.sh) * 16);
  }
  unsigned char* wsm = c1_smem + 2 * img_al;
  unsigned char* xch = wsm + (w_bytes + 1023) / 1024 * 1024;      // fused pool: two [Hp * ph rows][NT halves + 16 bytes] exchange buffers
  const uint32_t a_w = smem_u32(&bar_w), a_imgfull = smem_u32(&bar_imgfull[0]), a_imgempty = smem_u32(&bar_imgempty[0]);
  const uint32_t a_accfull = smem_u32(&bar_accfull[0]), a_accempty = smem_u32(&bar_accempty[0]);
  if (tid == 0) {
    mbar_init(a_w, 1);
    for (int k = 0; k < 2; ++k) {
      mbar_init(a_imgfull + 8 * k, 1);
      mbar_init(a_imgempty + 8 * k, C1_ISSUERS);
    }
    for (int k = 0; k < 4; ++k) {
      mbar_init(a_accfull + 8 * k, 1);
      mbar_init(a_accempty + 8 * k, C1_EPI);
    }
    fence_mbar_init();
  }
  for (int i = tid; i < p.NT; i += C1_THREADS) bias_s[i] = __ldg(p.bias + i);
  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;

  if (warp == 0) {
    // ======================================= producer =======================================
    if (lane == 0) {
      mbar_expect_tx(a_w, w_bytes);
      for (int off = 0; off < w_bytes; off += 32768) gt_bulk_load(smem_u32(wsm + off), p.B + off, min(32768, w_bytes - off), a_w);
      uint32_t n = 0;
      for (int clip = blockIdx.x; clip < p.n_clips; clip += gridDim.x, ++n) {
        const uint32_t ib = n & 1;
        if (n >= 2) mbar_wait(a_imgempty + 8 * ib, ((n >> 1) & 1) ^ 1);
        mbar_expect_tx(a_imgfull + 8 * ib, p.img_bytes);
        const unsigned char* src = p.X + static_cast<long long>(clip) * p.img_bytes;
        for (int off = 0; off < p.img_bytes; off += 32768)
          gt_bulk_load(smem_u32(c1_smem + ib * img_al + off), src + off, min(32768, p.img_bytes - off), a_imgfull + 8 * ib);
      }
    }
  } else if (warp < C1_EPI0) {
    // ======================================= MMA issuers ====================================
    const uint32_t me = warp - 1;                                  // this warp's accumulator = the tiles it issues (c & 1 == me)
    const uint32_t idesc = make_idesc(128, p.NT, 0);
    const uint64_t a_hi = make_desc(0, p.lbo, 128) & 0xFFFFFFFFFFFF0000ull;
    const uint64_t b_hi = make_desc(0, p.NT * 16, 128) & 0xFFFFFFFFFFFF0000ull;
    const uint32_t wb = smem_u32(wsm);
    uint32_t aoff_r[C1_UNROLL];
#pragma unroll
    for (int j = 0; j < C1_UNROLL; ++j) aoff_r[j] = c1_aoff[min(j, C1_MAXSTEPS - 1)] >> 4;
    mbar_wait(a_w, 0);
    uint32_t n = 0, c = 0;
    for (int clip = blockIdx.x; clip < p.n_clips; clip += gridDim.x, ++n) {
      const uint32_t ib = n & 1;
      mbar_wait(a_imgfull + 8 * ib, (n >> 1) & 1);
      fence_after();
      const uint32_t img = smem_u32(c1_smem + ib * img_al);
      for (int t = 0; t < p.n_tiles; ++t, ++c) {
        const uint32_t ab = c & acc_mask;
        if (C1_ISSUERS == 2 && (c & 1) != me) continue;
        if (c > acc_mask) {
          mbar_wait(a_accempty + 8 * ab, ((c >> acc_shift) & 1) ^ 1);
          fence_after();
        }
        if (gt_elect_one()) {
          const uint32_t d = tmem + ab * acc_stride;
          const uint32_t a_t = (img + t * 2048) >> 4, n_j = (p.mode & 4) ? 0 : p.KP / 2;
#pragma unroll
          for (uint32_t j = 0; j < C1_UNROLL; ++j)                     // offsets in registers: no load in front of an MMA
            if (j < n_j) mma_f16_ss(d, a_hi | (a_t + aoff_r[j]), b_hi | ((wb >> 4) + j * 2 * p.NT), idesc, j > 0 ? 1u : 0u);
          for (uint32_t j = C1_UNROLL; j < n_j; ++j)
            mma_f16_ss(d, a_hi | (a_t + (c1_aoff[j] >> 4)), b_hi | ((wb >> 4) + j * 2 * p.NT), idesc, 1u);
          commit(a_accfull + 8 * ab);
          if (t + C1_ISSUERS >= p.n_tiles) commit(a_imgempty + 8 * ib);   // this warp's last tile of the clip (n_tiles >= 2)
        }
        __syncwarp();
      }
    }
  } else {
    // ======================================= epilogue =======================================
    const int quad = warp & 3, row = quad * 32 + lane, part = (warp - C1_EPI0) >> 2;          // column part of this warp
    const uint32_t tlane = static_cast<uint32_t>(quad * 32) << 16;
    const int n16 = p.NT / 16, c_lo = part * n16 / (C1_EPI / 4), c_hi = (part + 1) * n16 / (C1_EPI / 4);   // 16-column groups
    uint32_t c = 0;
    if (p.fuse == 1) {
      // max over the pw coefficient columns in registers (tile = column, thread = time row), then bias + ReLU + f16 (all monotonic,
      // so the order against the reference's conv -> ReLU -> pool does not change a bit), max over ph time rows through shared memory
      constexpr int MAXG = (C1_FUSE_MAXNT / 16 + C1_EPI / 4 - 1) / (C1_EPI / 4);
      float rm[MAXG][16];
      const int n_g = c_hi - c_lo, xrow = p.NT * 2 + 16, etid = tid - C1_EPI0 * 32, rows_p = p.Hp * p.ph, upc = p.Hp * p.cpt;
      const int hp0 = etid / p.cpt, cg0 = etid - hp0 * p.cpt, d_hp = C1_EPI * 32 / p.cpt, d_cg = C1_EPI * 32 - d_hp * p.cpt;   // this thread's pooled units
      uint32_t xb = 0;
      for (int clip = blockIdx.x; clip < p.n_clips; clip += gridDim.x) {
        for (int t = 0, wq = 0, wp = 0; t < p.n_tiles; ++t, ++c) {
          const uint32_t ab = c & acc_mask;
          mbar_wait(a_accfull + 8 * ab, (c >> acc_shift) & 1);
          fence_after();
#pragma unroll
          for (int gi = 0; gi < MAXG; ++gi) {
            if (gi >= n_g) break;
            uint32_t v[16];
            tmem_ld16(tmem + tlane + ab * acc_stride + (c_lo + gi) * 16, v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 16; ++k) rm[gi][k] = wq == 0 ? __uint_as_float(v[k]) : fmaxf(rm[gi][k], __uint_as_float(v[k]));
          }
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(a_accempty + 8 * ab);
          if (++wq < p.pw) continue;
          wq = 0;
          if (wp >= p.Wp) continue;                                  // floor mode: the last Wo % pw columns are dropped
          unsigned char* xs = xch + xb * (rows_p * xrow);
          if (row < rows_p && !(p.mode & 2)) {
#pragma unroll
            for (int gi = 0; gi < MAXG; ++gi) {
              if (gi >= n_g) break;
              const int c0 = (c_lo + gi) * 16;
              float f[16];
#pragma unroll
              for (int k = 0; k < 16; k += 4) {
                const float4 b4 = *reinterpret_cast<const float4*>(&bias_s[c0 + k]);
                f[k] = fmaxf(rm[gi][k] + b4.x, 0.f);
                f[k + 1] = fmaxf(rm[gi][k + 1] + b4.y, 0.f);
                f[k + 2] = fmaxf(rm[gi][k + 2] + b4.z, 0.f);
                f[k + 3] = fmaxf(rm[gi][k + 3] + b4.w, 0.f);
              }
              uint4* d = reinterpret_cast<uint4*>(xs + row * xrow + c0 * 2);
              d[0] = make_uint4(gt_pack_h2(f[0], f[1]), gt_pack_h2(f[2], f[3]), gt_pack_h2(f[4], f[5]), gt_pack_h2(f[6], f[7]));
              d[1] = make_uint4(gt_pack_h2(f[8], f[9]), gt_pack_h2(f[10], f[11]), gt_pack_h2(f[12], f[13]), gt_pack_h2(f[14], f[15]));
            }
          }
          asm volatile("bar.sync 1, %0;" ::"n"(C1_EPI * 32) : "memory");             // the 8 epilogue warps: the column block is staged
          uint4* P = reinterpret_cast<uint4*>(p.P) + static_cast<long long>(clip) * p.cpt * p.R_alloc;
          for (int u = etid, hp = hp0, cg = cg0; u < upc && !(p.mode & 2); u += C1_EPI * 32) {
            const unsigned char* src = xs + (hp * p.ph) * xrow + cg * 16;
            uint4 o = *reinterpret_cast<const uint4*>(src);
            for (int i = 1; i < p.ph; ++i) {
              const uint4 q = *reinterpret_cast<const uint4*>(src + i * xrow);
              o.x = pk_hmax2(o.x, q.x); o.y = pk_hmax2(o.y, q.y); o.z = pk_hmax2(o.z, q.z); o.w = pk_hmax2(o.w, q.w);
            }
            P[static_cast<long long>(cg) * p.R_alloc + hp * p.Wp + wp] = o;
            hp += d_hp;                                              // the unit one stride further on, without a division
            cg += d_cg;
            if (cg >= p.cpt) {
              cg -= p.cpt;
              ++hp;
            }
          }
          xb ^= 1;                                                   // the other buffer is two barriers away from its readers
          ++wp;
        }
      }
    } else if (p.fuse == 2) {
      const int xrow = p.NT * 2 + 16, etid = tid - C1_EPI0 * 32, rows_p = p.Hp * p.ph, col_bytes = rows_p * xrow, upc = p.Hp * p.cpt;
      const int col_in_tile = row >> p.hc_shift, ho = row & ((1 << p.hc_shift) - 1);      // this thread's place in every tile
      const int hp0 = etid / p.cpt, cg0 = etid - hp0 * p.cpt, d_hp = C1_EPI * 32 / p.cpt, d_cg = C1_EPI * 32 - d_hp * p.cpt;   // its pooled units
      for (int clip = blockIdx.x; clip < p.n_clips; clip += gridDim.x) {
        uint4* P = reinterpret_cast<uint4*>(p.P) + static_cast<long long>(clip) * p.cpt * p.R_alloc;
        for (int t = 0, wp_done = 0, pslot = 0, tslot = 0; t < p.n_tiles; ++t, ++c) {   // ring slots of the next column block / tile
          const uint32_t ab = c & acc_mask;
          mbar_wait(a_accfull + 8 * ab, (c >> acc_shift) & 1);
          fence_after();
          const int col = (t << (7 - p.hc_shift)) + col_in_tile;
          const bool valid = ho < rows_p && col < p.Wp * p.pw && !(p.mode & 2);
          int slot = tslot + col_in_tile;
          if (slot >= p.n_cols) slot -= p.n_cols;
          unsigned char* xs = xch + slot * col_bytes + ho * xrow;
          tslot += 1 << (7 - p.hc_shift);
          if (tslot >= p.n_cols) tslot -= p.n_cols;
          if (c_lo == c_hi && lane == 0) mbar_arrive(a_accempty + 8 * ab);       // fewer column groups than parts
          for (int g = c_lo; g < c_hi; ++g) {
            const int c0 = g * 16;
            uint32_t v[16];
            tmem_ld16(tmem + tlane + ab * acc_stride + c0, v);
            tmem_ld_wait();
            if (g == c_hi - 1) {
              fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(a_accempty + 8 * ab);
            }
            if (!valid) continue;
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; k += 4) {
              const float4 b4 = *reinterpret_cast<const float4*>(&bias_s[c0 + k]);
              f[k] = fmaxf(__uint_as_float(v[k]) + b4.x, 0.f);
              f[k + 1] = fmaxf(__uint_as_float(v[k + 1]) + b4.y, 0.f);
              f[k + 2] = fmaxf(__uint_as_float(v[k + 2]) + b4.z, 0.f);
              f[k + 3] = fmaxf(__uint_as_float(v[k + 3]) + b4.w, 0.f);
            }
            uint4* d = reinterpret_cast<uint4*>(xs + c0 * 2);
            d[0] = make_uint4(gt_pack_h2(f[0], f[1]), gt_pack_h2(f[2], f[3]), gt_pack_h2(f[4], f[5]), gt_pack_h2(f[6], f[7]));
            d[1] = make_uint4(gt_pack_h2(f[8], f[9]), gt_pack_h2(f[10], f[11]), gt_pack_h2(f[12], f[13]), gt_pack_h2(f[14], f[15]));
          }
          asm volatile("bar.sync 1, %0;" ::"n"(C1_EPI * 32) : "memory");             // the 8 epilogue warps: the tile's columns are staged
          const int wp_end = min(p.Wp, min(p.Wo, ((t + 1) * 128) >> p.hc_shift) / p.pw);
          for (int wp = wp_done; wp < wp_end && !(p.mode & 2); ++wp) {
            for (int u = etid, hp = hp0, cg = cg0; u < upc; u += C1_EPI * 32) {
              const unsigned char* base = xch + (hp * p.ph) * xrow + cg * 16;
              uint4 o = make_uint4(0, 0, 0, 0);                      // the staged values are ReLU-ed: 0 is the identity of the max
              for (int j = 0, sj = pslot; j < p.pw; ++j) {
                const unsigned char* src = base + sj * col_bytes;
                for (int i = 0; i < p.ph; ++i) {
                  const uint4 q = *reinterpret_cast<const uint4*>(src + i * xrow);
                  o.x = pk_hmax2(o.x, q.x); o.y = pk_hmax2(o.y, q.y); o.z = pk_hmax2(o.z, q.z); o.w = pk_hmax2(o.w, q.w);
                }
                if (++sj == p.n_cols) sj = 0;
              }
              P[static_cast<long long>(cg) * p.R_alloc + hp * p.Wp + wp] = o;
              hp += d_hp;                                            // the unit one stride further on, without a division
              cg += d_cg;
              if (cg >= p.cpt) {
                cg -= p.cpt;
                ++hp;
              }
            }
            pslot += p.pw;
            if (pslot >= p.n_cols) pslot -= p.n_cols;
          }
          wp_done = wp_end;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(C1_EPI * 32) : "memory");               // the ring restarts at column 0 with the next clip
      }
    } else
    for (int clip = blockIdx.x; clip < p.n_clips; clip += gridDim.x) {
      for (int t = 0; t < p.n_tiles; ++t, ++c) {
        const uint32_t ab = c & acc_mask;
        mbar_wait(a_accfull + 8 * ab, (c >> acc_shift) & 1);
        fence_after();
        const int r = t * 128 + row, wo = r / p.Hph, ho = r - wo * p.Hph;
        const bool valid = ho < p.Ho && wo < p.Wo;
        __half* dst = p.C + ((static_cast<long long>(clip) * p.Ho + ho) * p.Wo + wo) * p.ldc;
        if (c_lo == c_hi && lane == 0) mbar_arrive(a_accempty + 8 * ab);         // fewer column groups than parts
        for (int g = c_lo; g < c_hi; ++g) {
          const int c0 = g * 16;
          uint32_t v[16];
          tmem_ld16(tmem + tlane + ab * acc_stride + c0, v);
          tmem_ld_wait();
          if (g == c_hi - 1) {                                       // this warp's share of the accumulator is in registers
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_accempty + 8 * ab);
          }
          if (!valid || (p.mode & 2)) continue;
          float f[16];
#pragma unroll
          for (int k = 0; k < 16; k += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(&bias_s[c0 + k]);
            f[k] = fmaxf(__uint_as_float(v[k]) + b4.x, 0.f);
            f[k + 1] = fmaxf(__uint_as_float(v[k + 1]) + b4.y, 0.f);
            f[k + 2] = fmaxf(__uint_as_float(v[k + 2]) + b4.z, 0.f);
            f[k + 3] = fmaxf(__uint_as_float(v[k + 3]) + b4.w, 0.f);
          }
          *reinterpret_cast<uint4*>(dst + c0) = make_uint4(gt_pack_h2(f[0], f[1]), gt_pack_h2(f[2], f[3]), gt_pack_h2(f[4], f[5]), gt_pack_h2(f[6], f[7]));
          *reinterpret_cast<uint4*>(dst + c0 + 8) =
              make_uint4(gt_pack_h2(f[8], f[9]), gt_pack_h2(f[10], f[11]), gt_pack_h2(f[12], f[13]), gt_pack_h2(f[14], f[15]));
        }
      }
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// A operand of one layer (see the file comment).  One thread per 16-byte unit (row m, K plane), consecutive threads =
// consecutive rows of one plane (coalesced 16-byte stores).  The previous layer's MaxPool(ph, pw) (floor mode, model.py:190,193)
// is applied while reading: max over the window of the stored, already ReLU-ed f16 activation (max commutes with the f16 rounding).
struct PackParams {
  const void* in;              // float32 [B][Hin][Win] (in_f32) or f16 NHWC [B][Hin][Win][Cs]
  unsigned char* A;
  long long M, n_units;        // rows = B * Ho * Wo; units = n_mtiles * KP * 128
  int in_f32, Hin, Win, Cs;    // stored input map (before the pool); Cs = channel stride (multiple of 8; 1 for float32)
  int ph, pw;                  // pool window applied to the input
  int KH, KW, sh, sw, Ho, Wo;  // kernel over the POOLED map, strides, output map
  int K, KP;                   // real K = KH * KW * Cs and planes (multiple of 2)
};

// One thread per ROW m of A: (clip, ho, wo) is decoded once, then the thread walks the K planes of its row.  Consecutive threads
// are consecutive rows, so every plane step of a warp is 32 adjacent 16-byte stores, and neighbouring rows read overlapping
// input (L1 hits).  The input of a pooled layer has been pooled by gemm_pool_kernel before (ph = pw = 1 here).
__global__ void __launch_bounds__(256) gemm_pack_kernel(const PackParams p) {
  const long long n_rows = (p.M + 127) / 128 * 128;
  for (long long m = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; m < n_rows; m += static_cast<long long>(gridDim.x) * 256) {
    uint4* dst = reinterpret_cast<uint4*>(p.A) + (m >> 7) * (static_cast<long long>(p.KP) * 128) + (m & 127);
    if (m >= p.M) {
      for (int pl = 0; pl < p.KP; ++pl) dst[static_cast<long long>(pl) * 128] = make_uint4(0, 0, 0, 0);
      continue;
    }
    const int P = p.Ho * p.Wo;
    const long long clip = m / P;
    const int pos = static_cast<int>(m - clip * P), ho = pos / p.Wo, wo = pos - ho * p.Wo;
    if (p.in_f32) {
      const float* x = static_cast<const float*>(p.in) + (clip * p.Hin + ho * p.sh) * p.Win + wo * p.sw;
      if (p.KW == 8) {                                       // every reference config: one kernel row = one plane
        for (int kh = 0; kh < p.KP; ++kh) {
          uint4 o = make_uint4(0, 0, 0, 0);
          if (kh < p.KH) {
            uint32_t h[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) h[e] = __half_as_ushort(__float2half_rn(fminf(fmaxf(__ldg(x + kh * p.Win + e), -60000.f), 60000.f)));
            o = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
          }
          dst[static_cast<long long>(kh) * 128] = o;
        }
      } else {
        for (int pl = 0; pl < p.KP; ++pl) {
          uint32_t h[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int k = pl * 8 + e;
            float v = 0.f;
            if (k < p.K) {
              const int kh = k / p.KW, kw = k - kh * p.KW;
              v = fminf(fmaxf(__ldg(x + kh * p.Win + kw), -60000.f), 60000.f);
            }
            h[e] = __half_as_ushort(__float2half_rn(v));
          }
          dst[static_cast<long long>(pl) * 128] = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
        }
      }
    } else {
      const int cpt = p.Cs >> 3;                             // planes per tap
      const uint4* base = reinterpret_cast<const uint4*>(static_cast<const __half*>(p.in) + ((clip * p.Hin + ho * p.sh) * p.Win + wo * p.sw) * p.Cs);
      int pl = 0;
      for (int kh = 0; kh < p.KH; ++kh)
        for (int kw = 0; kw < p.KW; ++kw) {
          const uint4* src = base + (kh * p.Win + kw) * cpt;
          for (int c = 0; c < cpt; ++c, ++pl) dst[static_cast<long long>(pl) * 128] = __ldg(src + c);
        }
      for (; pl < p.KP; ++pl) dst[static_cast<long long>(pl) * 128] = make_uint4(0, 0, 0, 0);
    }
  }
}

// The same operand with one thread per 16-byte unit (row, plane): for layers with few rows and a long K (the first linear layer
// after the convolutions: M = clips, K = the whole map), where a thread per row would leave most of the GPU idle.  NHWC input only.
__global__ void __launch_bounds__(256) gemm_pack_units_kernel(const PackParams p) {
  const int cpt = p.Cs >> 3, n_taps = p.KH * p.KW;
  for (long long u = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; u < p.n_units; u += static_cast<long long>(gridDim.x) * 256) {
    const int r = static_cast<int>(u & 127);
    const long long tp = u >> 7;
    const int plane = static_cast<int>(tp % p.KP);
    const long long m = (tp / p.KP) * 128 + r;
    uint4 o = make_uint4(0, 0, 0, 0);
    const int tap = plane / cpt, c = plane - tap * cpt;
    if (m < p.M && tap < n_taps) {
      const int P = p.Ho * p.Wo;
      const long long clip = m / P;
      const int pos = static_cast<int>(m - clip * P), ho = pos / p.Wo, wo = pos - ho * p.Wo;
      const int kh = tap / p.KW, kw = tap - kh * p.KW;
      o = __ldg(reinterpret_cast<const uint4*>(static_cast<const __half*>(p.in) + ((clip * p.Hin + ho * p.sh + kh) * p.Win + wo * p.sw + kw) * p.Cs) + c);
    }
    reinterpret_cast<uint4*>(p.A)[u] = o;
  }
}

// MaxPool(ph, pw), floor mode (model.py:190, 193), over an NHWC f16 map: [B][Hin][Win][Cs] -> [B][Hin / ph][Win / pw][Cs].
// One thread per 16-byte unit (8 channels) of the output.
__global__ void __launch_bounds__(256) gemm_pool_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, long long n_units, int Hin, int Win,
                                                        int cpt, int ph, int pw) {
  const int Hp = Hin / ph, Wp = Win / pw;
  for (long long u = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; u < n_units; u += static_cast<long long>(gridDim.x) * 256) {
    const int c = static_cast<int>(u % cpt);
    long long t = u / cpt;
    const int wp = static_cast<int>(t % Wp);
    t /= Wp;
    const int hp = static_cast<int>(t % Hp);
    const long long clip = t / Hp;
    const uint4* src = in + ((clip * Hin + hp * ph) * Win + wp * pw) * cpt + c;
    uint4 o = __ldg(src);
    for (int i = 0; i < ph; ++i)
      for (int j = 0; j < pw; ++j) {
        if (i == 0 && j == 0) continue;
        const uint4 q = __ldg(src + (i * Win + j) * cpt);
        o.x = pk_hmax2(o.x, q.x); o.y = pk_hmax2(o.y, q.y); o.z = pk_hmax2(o.z, q.z); o.w = pk_hmax2(o.w, q.w);
      }
    out[u] = o;
  }
}

// The same MaxPool, written as the PLANAR image the implicit-mode GEMM reads: out[clip][cg][R_alloc rows][8 halves], row =
// hp * Wp + wp; rows past the map (the tiles' and taps' over-read) are zero.  One thread per 16-byte unit of the output.
__global__ void __launch_bounds__(256) gemm_pool_planar_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, long long n_units, int Hin,
                                                               int Win, int cpt, int ph, int pw, int R_alloc) {
  const int Hp = Hin / ph, Wp = Win / pw;
  for (long long u = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; u < n_units; u += static_cast<long long>(gridDim.x) * 256) {
    const int r = static_cast<int>(u % R_alloc);
    long long t = u / R_alloc;
    const int c = static_cast<int>(t % cpt);
    const long long clip = t / cpt;
    uint4 o = make_uint4(0, 0, 0, 0);
    if (r < Hp * Wp) {
      const int hp = r / Wp, wp = r - hp * Wp;
      const uint4* src = in + ((clip * Hin + hp * ph) * Win + wp * pw) * cpt + c;
      o = __ldg(src);
      for (int i = 0; i < ph; ++i)
        for (int j = 0; j < pw; ++j) {
          if (i == 0 && j == 0) continue;
          const uint4 q = __ldg(src + (i * Win + j) * cpt);
          o.x = pk_hmax2(o.x, q.x); o.y = pk_hmax2(o.y, q.y); o.z = pk_hmax2(o.z, q.z); o.w = pk_hmax2(o.w, q.w);
        }
    }
    out[u] = o;
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct GemmLayer {
  // geometry
  int in_f32 = 0, Hin = 0, Win = 0, Cs = 1, ph = 1, pw = 1, KH = 1, KW = 1, sh = 1, sw = 1, Ho = 1, Wo = 1;
  int K = 0, KP = 0, N = 0, NT = 16, n_ntiles = 1, relu = 0;
  int imp = 0, R_alloc = 0, tiles_per_clip = 0;     // implicit mode (stride-1 conv over the pooled planar map): no A operand
  int res = 0, res_smem = 0, res_stages = 4, res_G = 1;   // ... with the clips' images resident in shared memory (conv_res_kernel)
  int c1res = 0, c1_units = 0, c1_tiles = 0, c1_hph = 0, c1_smem = 0;   // conv1 over the resident x8p image (conv1_res_kernel)
  int c1_fuse = 0, c1_cols = 0, c1_shift = 0;   // ... whose epilogue applies the MaxPool and writes the next layer's planar image (1: columns in registers, 2: staged)
  int planar_in = 0;           // the previous layer wrote this layer's planar image (no gemm_pool_planar_kernel pass)
  unsigned char* d_B = nullptr;
  float* d_bias = nullptr;
  int ldc() const { return NT * n_ntiles; }
};

struct CnnGemmPlan {
  std::vector<GemmLayer> layers;
  int n_labels = 12, n_sms = 148, H0 = 101, W0 = 40;
  int64_t chunk = 1024;                 // clips per pass (bounds the workspace: the im2col A operand is the large buffer)
  int64_t a_bytes_per_clip = 0, act_bytes_per_clip = 0, pool_bytes_per_clip = 0;
  std::vector<void*> owned;
};

void cnn_gemm_plan_free(CnnGemmPlan* p) {
  if (!p) return;
  for (void* q : p->owned) cudaFree(q);
  delete p;
}

namespace {

template <typename T>
int gt_up(CnnGemmPlan* p, const std::vector<T>& h, T** d) {
  void* ptr = nullptr;
  HONK_CUDA(cudaMalloc(&ptr, std::max<size_t>(h.size(), 1) * sizeof(T)));
  p->owned.push_back(ptr);
  HONK_CUDA(cudaMemcpy(ptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = static_cast<T*>(ptr);
  return HONK_OK;
}

// Packs one layer's weights: wk(n, kh, kw, c) is the weight of output n for tap (kh, kw) and REAL input channel c < Cin;
// stored input channels c >= Cin (the previous layer's padding) get zero.
template <typename F>
int gt_pack_weights(CnnGemmPlan* p, GemmLayer& L, int Cin, const std::vector<float>& bias, F wk) {
  L.K = L.KH * L.KW * L.Cs;
  L.KP = (L.K + 15) / 16 * 2;
  const int npad = (L.N + 15) / 16 * 16;
  L.n_ntiles = (npad + GT_MAXNT - 1) / GT_MAXNT;
  L.NT = ((npad + L.n_ntiles - 1) / L.n_ntiles + 15) / 16 * 16;
  std::vector<__half> B(static_cast<size_t>(L.n_ntiles) * L.KP * L.NT * 8, __float2half(0.f));
  std::vector<float> b(static_cast<size_t>(L.n_ntiles) * L.NT, 0.f);
  for (int n = 0; n < L.N; ++n) {
    const int nt = n / L.NT, nr = n - nt * L.NT;
    b[n] = bias[n];
    for (int kh = 0; kh < L.KH; ++kh)
      for (int kw = 0; kw < L.KW; ++kw)
        for (int c = 0; c < Cin; ++c) {
          const int k = (kh * L.KW + kw) * L.Cs + c;
          B[((static_cast<size_t>(nt) * L.KP + k / 8) * L.NT + nr) * 8 + k % 8] = __float2half(wk(n, kh, kw, c));
        }
  }
  __half* d = nullptr;
  HONK_TRY(gt_up(p, B, &d));
  L.d_B = reinterpret_cast<unsigned char*>(d);
  HONK_TRY(gt_up(p, b, &L.d_bias));
  return HONK_OK;
}

}  // namespace

// Every cnn-* configuration (model.py:385-407); *out stays nullptr only when a layer does not fit the kernel's limits.
int cnn_gemm_plan_create(honk_model* m, CnnGemmPlan** out) {
  *out = nullptr;
  const honk_model_config& c = m->cfg;
  if (c.family != HONK_FAMILY_CNN) return HONK_OK;
  if (const char* e = getenv("HONK_DISABLE_GEMM_TC"))
    if (atoi(e)) return HONK_OK;
  std::unique_ptr<CnnGemmPlan, void (*)(CnnGemmPlan*)> p(new CnnGemmPlan(), cnn_gemm_plan_free);
  p->n_labels = c.n_labels;
  p->n_sms = m->n_sms;
  p->H0 = m->H0;
  p->W0 = m->W0;
  auto conv_out = [](int n, int k, int s) { return (n - k) / s + 1; };
  // ---- conv1 (model.py:187-190): float32 MFCC image in, NHWC f16 out
  int H, W, C, Cs, ph, pw;                      // stored map of the latest conv layer and the pool still to be applied to it
  {
    GemmLayer L;
    L.in_f32 = 1; L.Hin = m->H0; L.Win = m->W0; L.Cs = 1;
    L.KH = c.conv1_size[0]; L.KW = c.conv1_size[1]; L.sh = c.conv1_stride[0]; L.sw = c.conv1_stride[1];
    L.Ho = conv_out(m->H0, L.KH, L.sh); L.Wo = conv_out(m->W0, L.KW, L.sw);
    L.N = c.n_feature_maps1; L.relu = 1;
    if (L.Ho < 1 || L.Wo < 1) return HONK_OK;
    const HostTensor& w = m->host.at("conv1.weight");
    const int KH = L.KH, KW = L.KW;
    HONK_TRY(gt_pack_weights(p.get(), L, 1, m->host.at("conv1.bias").data,
                             [&](int n, int kh, int kw, int) { return w.data[(static_cast<size_t>(n) * KH + kh) * KW + kw]; }));
    {
      // conv1 without an im2col operand when it is a real convolution in time (KW = 8 coefficient taps, coefficient stride 1)
      static const bool no_c1 = getenv("HONK_GEMM_CONV1_RESIDENT") && atoi(getenv("HONK_GEMM_CONV1_RESIDENT")) == 0;
      int max_smem_ = 0;
      HONK_CUDA(cudaDeviceGetAttribute(&max_smem_, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
      const int sh = L.sh, Wx = m->W0 - 7, Hph = (m->H0 + sh - 1) / sh;
      // time strides 4 and 8 (cnn-tstride4/8): too few rows per phase image to beat the im2col operand on its own (measured); kept
      // only if the epilogue can also take over the MaxPool (decided below, once conv2's operand is known)
      static const int c1_maxsh = getenv("HONK_C1_MAXSH") ? atoi(getenv("HONK_C1_MAXSH")) : 8;
      if (!no_c1 && L.KW == 8 && L.sw == 1 && L.n_ntiles == 1 && (sh == 1 || (sh % 2 == 0 && sh <= c1_maxsh)) && L.Ho * L.Wo >= 256 && L.Wo == Wx) {
        L.c1_hph = Hph;
        L.c1_tiles = (Wx * Hph + 127) / 128;
        L.c1_units = ((sh - 1) * Wx * Hph + L.c1_tiles * 128 + L.KH / sh + 2 + 7) / 8 * 8;
        L.c1_units = std::max(L.c1_units, sh * Wx * Hph);
        const int img_al = (L.c1_units * 16 + 1023) / 1024 * 1024;
        L.c1_smem = 2 * img_al + L.KP * L.NT * 16 + 1024;
        L.c1res = L.c1_smem + 2048 <= max_smem_ && L.KP / 2 <= C1_MAXSTEPS && (sh == 1 ? 16 : Wx * Hph * 16) < (1 << 18);
      }
    }
    H = L.Ho; W = L.Wo; C = L.N; Cs = L.ldc(); ph = c.conv1_pool[0]; pw = c.conv1_pool[1];
    p->layers.push_back(L);
  }
  // ---- conv2 (model.py:191-194)
  if (c.has_conv2) {
    GemmLayer L;
    L.Hin = H; L.Win = W; L.Cs = Cs; L.ph = ph; L.pw = pw;
    L.KH = c.conv2_size[0]; L.KW = c.conv2_size[1]; L.sh = c.conv2_stride[0]; L.sw = c.conv2_stride[1];
    L.Ho = conv_out(H / ph, L.KH, L.sh); L.Wo = conv_out(W / pw, L.KW, L.sw);
    L.N = c.n_feature_maps2; L.relu = 1;
    if (L.Ho < 1 || L.Wo < 1) return HONK_OK;
    const HostTensor& w = m->host.at("conv2.weight");
    const int KH = L.KH, KW = L.KW, Cin = C;
    HONK_TRY(gt_pack_weights(p.get(), L, Cin, m->host.at("conv2.bias").data, [&](int n, int kh, int kw, int ci) {
      return w.data[((static_cast<size_t>(n) * Cin + ci) * KH + kh) * KW + kw];
    }));
    // implicit mode: a stride-1 convolution whose input went through a MaxPool reads the pooled map as a planar image by
    // tap offsets instead of a materialised im2col operand (HONK_GEMM_IMPLICIT=0: the im2col path)
    static const bool no_imp = getenv("HONK_GEMM_IMPLICIT") && atoi(getenv("HONK_GEMM_IMPLICIT")) == 0;
    if (!no_imp && L.sh == 1 && L.sw == 1 && L.ph * L.pw > 1 && L.Ho * L.Wo >= 32) {
      const int Wp = L.Win / L.pw;
      L.imp = 1;
      L.tiles_per_clip = ((L.Ho - 1) * Wp + L.Wo + 127) / 128;
      L.R_alloc = (L.tiles_per_clip * 128 + (L.KH - 1) * Wp + L.KW + 7) / 8 * 8;
      int max_smem_ = 0;
      HONK_CUDA(cudaDeviceGetAttribute(&max_smem_, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
      const int img1 = (L.Cs >> 3) * L.R_alloc * 16, b_stage = GT_KC * L.NT * 16;
      // clips per pass: as many as TMEM (512 columns) and the shared memory next to a 4-stage weight ring hold, at most 4
      L.res_G = std::max(1, std::min({4, 512 / std::max(1, L.tiles_per_clip * L.NT), (max_smem_ - 6144 - 4 * b_stage) / std::max(1, img1)}));
      if (const char* e = getenv("HONK_CR_G")) L.res_G = std::max(1, std::min(L.res_G, atoi(e)));
      const int img = (L.res_G * img1 + 1023) / 1024 * 1024;
      L.res_stages = std::max(2, std::min(CR_MAXSTAGE, (max_smem_ - 6144 - img) / b_stage));
      if (const char* e = getenv("HONK_CR_STAGES")) L.res_stages = std::max(2, std::min(L.res_stages, atoi(e)));
      L.res_smem = img + L.res_stages * b_stage + 1024;
      static const bool no_res = getenv("HONK_GEMM_RESIDENT") && atoi(getenv("HONK_GEMM_RESIDENT")) == 0;
      L.res = !no_res && L.n_ntiles == 1 && L.tiles_per_clip * L.NT <= 512 && L.KP / 2 <= CR_MAXSTEPS && L.res_smem + 5120 <= max_smem_ && L.R_alloc < 16384;
      // measured (profiles/r04_configs.json): the resident kernel wins when a pass over the weight stream feeds at least two M
      // tiles (cnn-tpool2/3: three tiles of one clip; tstride2/4: two clips of three / two tiles; tstride8: three clips of one
      // tile; conv2 + its operand 3.1 -> 1.5, 3.3 -> 1.0, 1.5 -> 0.67, 0.66 -> 0.38 ms per 4096 clips); with a single tile per
      // pass every tile streams the whole weight tensor and the materialised operand is faster
      if (!L.res || L.tiles_per_clip * L.res_G < 2) L.imp = L.res = 0;
      // conv1's epilogue can do the MaxPool and write this image itself when a thread keeps one time row across the
      // coefficient columns (column stride 128 in the x8p image) and the column maxima fit its registers
      GemmLayer& L1 = p->layers[0];
      static const bool no_fuse = getenv("HONK_C1_FUSE") && atoi(getenv("HONK_C1_FUSE")) == 0;
      const int Wx = m->W0 - 7, w_al = (L1.KP * L1.NT * 16 + 1023) / 1024 * 1024, rows_p = (L.Hin / L.ph) * L.ph;
      if (!no_fuse && L.res && L1.c1res && L1.sh == 1 && m->H0 <= 128 && L1.NT <= C1_FUSE_MAXNT && L.pw <= L1.Wo && L.ph <= L1.Ho) {
        const int units = (Wx * 128 + L1.KH + 2 + 7) / 8 * 8, img_al = (units * 16 + 1023) / 1024 * 1024;
        const int smem = 2 * img_al + w_al + 2 * rows_p * (L1.NT * 2 + 16) + 1024;
        if (smem + 2048 <= max_smem_) {
          L1.c1_fuse = 1; L1.c1_hph = 128; L1.c1_tiles = Wx; L1.c1_units = units; L1.c1_smem = smem;
          L.planar_in = 1;
        }
      } else if (!no_fuse && L.res && L1.c1res && L1.sh > 1 && L1.c1_hph <= 64 && L.pw <= L1.Wo && L.ph <= L1.Ho) {
        int shift = 0;
        while ((1 << shift) < L1.c1_hph) ++shift;
        const int Hc = 1 << shift, tiles = (Wx * Hc + 127) / 128, n_cols = 2 * (128 / Hc) + L.pw - 1;
        const int units = ((L1.sh - 1) * Wx * Hc + tiles * 128 + L1.KH / L1.sh + 2 + 7) / 8 * 8, img_al = (units * 16 + 1023) / 1024 * 1024;
        const int smem = 2 * img_al + w_al + n_cols * rows_p * (L1.NT * 2 + 16) + 1024;
        if (smem + 2048 <= max_smem_ && Wx * Hc * 16 < (1 << 18)) {
          L1.c1_fuse = 2; L1.c1_hph = Hc; L1.c1_tiles = tiles; L1.c1_units = std::max(units, L1.sh * Wx * Hc); L1.c1_smem = smem;
          L1.c1_cols = n_cols; L1.c1_shift = shift;
          L.planar_in = 1;
          if (getenv("HONK_C1_FUSE") && atoi(getenv("HONK_C1_FUSE")) == 3) L1.c1_fuse = 0, L.planar_in = 0, L1.c1_smem = 2 * img_al + L1.KP * L1.NT * 16 + 1024;   // geometry only
        }
      }
    }
    H = L.Ho; W = L.Wo; C = L.N; Cs = L.ldc(); ph = c.conv2_pool[0]; pw = c.conv2_pool[1];
    p->layers.push_back(L);
  }
  if (p->layers[0].c1res && p->layers[0].sh > 2 && !p->layers[0].c1_fuse) p->layers[0].c1res = 0;   // see conv1 above
  // ---- linear layers (model.py:195-205): the first one is a convolution whose kernel covers the whole pooled map, with the
  // weights permuted from the reference's (C, H, W) flatten order; the others are 1 x 1 over a 1 x 1 map
  std::vector<std::pair<std::string, int>> fcs;       // (name, relu)
  if (!c.tf_variant) fcs.push_back({"lin", 0});
  if (c.dnn1_size > 0) {
    fcs.push_back({"dnn1", c.tf_variant ? 0 : 1});
    if (c.dnn2_size > 0) fcs.push_back({"dnn2", 0});
  }
  fcs.push_back({"output", 0});
  bool first = true;
  for (auto& fc : fcs) {
    const HostTensor& w = m->host.at(fc.first + ".weight");
    GemmLayer L;
    L.Hin = H; L.Win = W; L.Cs = Cs; L.ph = ph; L.pw = pw;
    L.KH = H / ph; L.KW = W / pw; L.Ho = 1; L.Wo = 1;
    L.N = static_cast<int>(w.dims[0]); L.relu = fc.second;
    const int KH = L.KH, KW = L.KW, Cin = C;
    if (KH < 1 || KW < 1 || static_cast<int64_t>(w.dims[1]) != static_cast<int64_t>(Cin) * KH * KW)
      return set_error(HONK_ERR_KEY, "size mismatch for \"%s.weight\"", fc.first.c_str());
    (void)first;
    HONK_TRY(gt_pack_weights(p.get(), L, Cin, m->host.at(fc.first + ".bias").data, [&](int n, int kh, int kw, int ci) {
      return w.data[static_cast<size_t>(n) * Cin * KH * KW + (static_cast<size_t>(ci) * KH + kh) * KW + kw];   // flatten index c*H*W + h*W + w
    }));
    H = 1; W = 1; C = L.N; Cs = L.ldc(); ph = 1; pw = 1;
    first = false;
    p->layers.push_back(L);
  }
  // workspace per clip: the largest A operand and the largest activation (two ping-pong buffers)
  for (const GemmLayer& L : p->layers) {
    if (L.c1res) p->a_bytes_per_clip = std::max<int64_t>(p->a_bytes_per_clip, static_cast<int64_t>(L.c1_units) * 16);
    else if (!L.imp) p->a_bytes_per_clip = std::max<int64_t>(p->a_bytes_per_clip, static_cast<int64_t>(L.Ho) * L.Wo * L.KP * 16);
    p->act_bytes_per_clip = std::max<int64_t>(p->act_bytes_per_clip, static_cast<int64_t>(L.Ho) * L.Wo * L.ldc() * 2);
    if (L.imp)
      p->pool_bytes_per_clip = std::max<int64_t>(p->pool_bytes_per_clip, static_cast<int64_t>(L.Cs >> 3) * L.R_alloc * 16);
    else if (!L.in_f32 && L.ph * L.pw > 1)
      p->pool_bytes_per_clip = std::max<int64_t>(p->pool_bytes_per_clip, static_cast<int64_t>(L.Hin / L.ph) * (L.Win / L.pw) * L.Cs * 2);
    if (L.KP > 65535 * 2) return HONK_OK;
  }
  // clips per pass: keep the A operand of a pass under ~1 GB
  p->chunk = std::max<int64_t>(128, std::min<int64_t>(4096, (1ll << 30) / std::max<int64_t>(p->a_bytes_per_clip, 1) / 128 * 128));
  if (const char* e = getenv("HONK_GEMM_CHUNK")) p->chunk = std::max(1, atoi(e));      // tests: several passes on a small batch
  int max_smem = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
  const int smem = GT_NSTAGE * (GT_KC * 2048 + GT_KC * GT_MAXNT * 16) + 1024;
  if (smem + 1024 > max_smem) return HONK_OK;
  HONK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  {
    int res_smem = 0;
    for (const GemmLayer& L : p->layers)
      if (L.res) res_smem = std::max(res_smem, L.res_smem);
    if (res_smem) {
      HONK_CUDA(cudaFuncSetAttribute(conv_res_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, res_smem));
      HONK_CUDA(cudaFuncSetAttribute(conv_res_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, res_smem));
      HONK_CUDA(cudaFuncSetAttribute(conv_res_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, res_smem));
      HONK_CUDA(cudaFuncSetAttribute(conv_res_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, res_smem));
      HONK_CUDA(cudaFuncSetAttribute(conv_res_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, res_smem));
    }
    if (p->layers[0].c1res) HONK_CUDA(cudaFuncSetAttribute(conv1_res_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, p->layers[0].c1_smem));
  }
  *out = p.release();
  return HONK_OK;
}

static int64_t gt_a_bytes(const CnnGemmPlan* p, int64_t n) {
  int64_t mx = 0;
  for (const GemmLayer& L : p->layers) {
    if (L.c1res) mx = std::max<int64_t>(mx, n * L.c1_units * 16);
    else if (!L.imp) mx = std::max<int64_t>(mx, (n * L.Ho * L.Wo + 127) / 128 * L.KP * 2048);
  }
  return mx;
}

int64_t cnn_gemm_workspace_bytes(const CnnGemmPlan* p, int64_t n_clips) {
  const int64_t n = std::min(n_clips, p->chunk);
  Arena a(nullptr, 0);
  a.take<unsigned char>(gt_a_bytes(p, n));
  a.take<unsigned char>(n * p->act_bytes_per_clip);
  a.take<unsigned char>(n * p->act_bytes_per_clip);
  a.take<unsigned char>(n * p->pool_bytes_per_clip);
  return align_up(a.off, 256);
}

int cnn_gemm_forward(const CnnGemmPlan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                     cudaStream_t s) {
  const int64_t nmax = std::min(n_clips, p->chunk);
  Arena ar(ws, ws_bytes);
  unsigned char* A = ar.take<unsigned char>(gt_a_bytes(p, nmax));
  unsigned char* act[2];
  act[0] = ar.take<unsigned char>(nmax * p->act_bytes_per_clip);
  act[1] = ar.take<unsigned char>(nmax * p->act_bytes_per_clip);
  unsigned char* pooled = ar.take<unsigned char>(nmax * p->pool_bytes_per_clip);
  HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: tcgen05 workspace overrun");
  ProfSpan span("cnn_conv", s);
  for (int64_t c0 = 0; c0 < n_clips; c0 += p->chunk) {
    const int64_t n = std::min(p->chunk, n_clips - c0);
    const void* in = d_mfcc + c0 * p->H0 * p->W0;
    for (size_t li = 0; li < p->layers.size(); ++li) {
      const GemmLayer& L = p->layers[li];
      const bool last = li + 1 == p->layers.size();
      static const char* kPack[] = {"gt_pack0", "gt_pack1", "gt_pack2", "gt_pack3", "gt_pack4", "gt_pack5"};
      static const char* kGemm[] = {"gt_gemm0", "gt_gemm1", "gt_gemm2", "gt_gemm3", "gt_gemm4", "gt_gemm5"};
      if (L.c1res) {
        {
          ProfSpan sp(kPack[0], s);
          static const bool pack_units = getenv("HONK_C1_PACK_UNITS") && atoi(getenv("HONK_C1_PACK_UNITS")) != 0;
          if (p->W0 == 40 && !pack_units)
            x8p_pack_rows40_kernel<<<static_cast<unsigned>(n), 128, 0, s>>>(static_cast<const float*>(in), A, p->H0, L.sh, L.c1_hph, L.c1_units);
          else
            x8p_pack_kernel<<<static_cast<unsigned>(n), 256, 0, s>>>(static_cast<const float*>(in), A, p->H0, p->W0, L.sh, L.c1_hph, L.c1_units);
          HONK_LAUNCH_CHECK();
        }
        Conv1ResParams r{};
        r.X = A; r.B = L.d_B; r.bias = L.d_bias; r.C = reinterpret_cast<__half*>(act[li & 1]);
        r.n_clips = static_cast<int>(n); r.img_bytes = L.c1_units * 16; r.KP = L.KP; r.NT = L.NT; r.ldc = L.ldc(); r.Hph = L.c1_hph;
        r.Ho = L.Ho; r.Wo = L.Wo; r.n_tiles = L.c1_tiles; r.sh = L.sh; r.phase_stride = (p->W0 - 7) * L.c1_hph;
        r.lbo = L.sh == 1 ? 16 : r.phase_stride * 16;
        if (L.c1_fuse) {
          const GemmLayer& L2 = p->layers[1];
          r.fuse = L.c1_fuse; r.hc_shift = L.c1_shift; r.n_cols = L.c1_cols; r.ph = L2.ph; r.pw = L2.pw; r.Hp = L2.Hin / L2.ph; r.Wp = L2.Win / L2.pw; r.R_alloc = L2.R_alloc; r.cpt = L2.Cs >> 3;
          r.P = pooled;
        }
        static const int c1_mode = getenv("HONK_C1_MODE") ? atoi(getenv("HONK_C1_MODE")) : 0;
        r.mode = c1_mode;
        {
          ProfSpan sp(kGemm[0], s);
          conv1_res_kernel<<<static_cast<unsigned>(std::min<long long>(n, p->n_sms)), C1_THREADS, L.c1_smem, s>>>(r);
          HONK_LAUNCH_CHECK();
        }
        in = act[li & 1];
        continue;
      }
      int Hin = L.Hin, Win = L.Win;
      PackParams pk{};
      pk.M = n * L.Ho * L.Wo;
      long long n_mtiles = (pk.M + 127) / 128;
      if (L.imp && L.planar_in) {
        n_mtiles = n * L.tiles_per_clip;                   // conv1's epilogue wrote the pooled planar image
      } else if (L.imp) {
        // MaxPool into the planar image the GEMM's producer reads by tap offsets: no A operand for this layer
        ProfSpan sp(kPack[std::min<size_t>(li, 5)], s);
        const long long units = n * (L.Cs >> 3) * L.R_alloc;
        gemm_pool_planar_kernel<<<static_cast<unsigned>(std::min<long long>((units + 255) / 256, 32ll * p->n_sms)), 256, 0, s>>>(
            static_cast<const uint4*>(in), reinterpret_cast<uint4*>(pooled), units, L.Hin, L.Win, L.Cs >> 3, L.ph, L.pw, L.R_alloc);
        HONK_LAUNCH_CHECK();
        n_mtiles = n * L.tiles_per_clip;
      } else {
        if (!L.in_f32 && L.ph * L.pw > 1) {                  // the previous layer's MaxPool, once, before the im2col reads it many times
          ProfSpan sp(kPack[std::min<size_t>(li, 5)], s);
          Hin = L.Hin / L.ph;
          Win = L.Win / L.pw;
          const long long units = n * Hin * Win * (L.Cs >> 3);
          gemm_pool_kernel<<<static_cast<unsigned>(std::min<long long>((units + 255) / 256, 32ll * p->n_sms)), 256, 0, s>>>(
              static_cast<const uint4*>(in), reinterpret_cast<uint4*>(pooled), units, L.Hin, L.Win, L.Cs >> 3, L.ph, L.pw);
          HONK_LAUNCH_CHECK();
          in = pooled;
        }
        pk.in = in; pk.A = A;
        pk.n_units = n_mtiles * L.KP * 128;
        pk.in_f32 = L.in_f32; pk.Hin = Hin; pk.Win = Win; pk.Cs = L.Cs; pk.ph = 1; pk.pw = 1;
        pk.KH = L.KH; pk.KW = L.KW; pk.sh = L.sh; pk.sw = L.sw; pk.Ho = L.Ho; pk.Wo = L.Wo; pk.K = L.K; pk.KP = L.KP;
        ProfSpan sp(kPack[std::min<size_t>(li, 5)], s);
        if (!L.in_f32 && n_mtiles * 128 < 512ll * p->n_sms && L.KP >= 64)
          gemm_pack_units_kernel<<<static_cast<unsigned>(std::min<long long>((pk.n_units + 255) / 256, 32ll * p->n_sms)), 256, 0, s>>>(pk);
        else
          gemm_pack_kernel<<<static_cast<unsigned>(std::min<long long>((n_mtiles * 128 + 255) / 256, 32ll * p->n_sms)), 256, 0, s>>>(pk);
        HONK_LAUNCH_CHECK();
      }
      GemmTcParams g{};
      g.A = A; g.B = L.d_B; g.bias = L.d_bias;
      g.C = last ? static_cast<void*>(d_logits + c0 * p->n_labels) : static_cast<void*>(act[li & 1]);
      g.M = pk.M; g.KP = L.KP; g.NT = L.NT; g.n_ntiles = L.n_ntiles;
      g.ldc = last ? p->n_labels : L.ldc();
      g.n_store = last ? p->n_labels : L.ldc();
      g.relu = L.relu; g.out_f32 = last ? 1 : 0;
      g.n_mtiles = n_mtiles;
      if (L.imp) {
        g.imp = 1; g.Ap = pooled; g.NCG = L.Cs >> 3; g.R_alloc = L.R_alloc; g.Wp = L.Win / L.pw; g.KW = L.KW; g.n_taps = L.KH * L.KW;
        g.tiles_per_clip = L.tiles_per_clip; g.Ho = L.Ho; g.Wo = L.Wo;
      }
      const long long n_tiles = n_mtiles * L.n_ntiles;
      const int smem = GT_NSTAGE * (GT_KC * 2048 + GT_KC * L.NT * 16) + 1024;
      if (L.res) {
        ConvResParams r{};
        r.Ap = pooled; r.B = L.d_B; r.bias = L.d_bias; r.C = reinterpret_cast<__half*>(act[li & 1]);
        r.n_clips = static_cast<int>(n); r.NCG = L.Cs >> 3; r.R_alloc = L.R_alloc; r.Wp = L.Win / L.pw; r.KW = L.KW; r.n_taps = L.KH * L.KW;
        r.tiles_per_clip = L.tiles_per_clip; r.Ho = L.Ho; r.Wo = L.Wo; r.NT = L.NT; r.ldc = L.ldc(); r.relu = L.relu; r.n_stage = L.res_stages; r.G = L.res_G;
        static const int cr_mode = getenv("HONK_CR_MODE") ? atoi(getenv("HONK_CR_MODE")) : 0;
        r.mode = cr_mode;
        ProfSpan sp(kGemm[std::min<size_t>(li, 5)], s);
        const unsigned grid = static_cast<unsigned>(std::min<long long>((n + L.res_G - 1) / L.res_G, p->n_sms));
        switch (L.tiles_per_clip) {
          case 1: conv_res_kernel<1><<<grid, CR_THREADS, L.res_smem, s>>>(r); break;
          case 2: conv_res_kernel<2><<<grid, CR_THREADS, L.res_smem, s>>>(r); break;
          case 3: conv_res_kernel<3><<<grid, CR_THREADS, L.res_smem, s>>>(r); break;
          case 4: conv_res_kernel<4><<<grid, CR_THREADS, L.res_smem, s>>>(r); break;
          default: conv_res_kernel<0><<<grid, CR_THREADS, L.res_smem, s>>>(r); break;
        }
        HONK_LAUNCH_CHECK();
      } else {
        ProfSpan sp(kGemm[std::min<size_t>(li, 5)], s);
        gemm_tc_kernel<<<static_cast<unsigned>(std::min<long long>(n_tiles, p->n_sms)), GT_THREADS, smem, s>>>(g);
        HONK_LAUNCH_CHECK();
      }
      in = act[li & 1];
    }
  }
  return HONK_OK;
}

}  // namespace honk
