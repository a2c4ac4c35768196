// Integer-limb DFT of the front end on the tcgen05 tensor cores (kind::i8), for int16 PCM and the default geometry
// (n_fft 480, hop 160, bins 1..119).  Replaces the FP64 DMMA contraction of fe2_mfcc_kernel for the STFT stage of
// AudioPreprocessor.compute_mfccs (utils/manage_audio.py:31-38: librosa melspectrogram's stft).
//
// Why integers.  The reference computes the STFT in float64 and log() turns the RELATIVE error of a mel energy into
// absolute MFCC error, so a reduced-precision float DFT (error proportional to the frame's energy, not the bin's) is not
// an option.  An int8 x int8 -> int32 tensor-core product is EXACT, and sm_100a's kind::i8 MMA runs at twice the f16
// rate (scripts/probes/i8_probe.cu, profiles/r04_i8_probe.txt).  So
//   x_n  (int16 PCM)             = 256 xh_n + xl_n           xh signed, xl unsigned 8-bit limbs         (exact)
//   B_nk = w_n cos / -sin(2 pi k n / N), |B| <= 1,  round(B 2^F) = sum_j b_j 256^(J-j)   J balanced signed digits
//   sum_n x_n round(B_nk 2^F) = sum_{i=1..J+1} a_i 256^(J+1-i),   a_1 = xh.b_1, a_i = xh.b_i + xl.b_(i-1), a_(J+1) = xl.b_J
// with every a_i an exact int32 accumulation in TMEM (|a_i| < 2^25) and the Horner sum over i done in FP64 by the
// epilogue.  The only approximation is the rounding of B to F = 46 fractional bits: |dX| <~ 0.3 * 2^-46 * ||x||_2,
// about 5x the rounding noise of the float64 FFT the reference uses and ~1e-9 of one PCM LSB - far below the float32
// (complex64) rounding the reference applies to the spectrum next (librosa 0.6 stft dtype).
//
// GEMM shape.  A frame is three hop-sized segments, so the A operand is the PCM itself: row u of a clip = samples
// [-240 + 160 u, -240 + 160 u + 160) (reflect-padded at the clip ends), frame t = rows t, t+1, t+2 - the segment shift is a
// start-address offset of the A descriptor (16 bytes per row in the K-major no-swizzle layout), exactly like the time taps
// of the conv kernels.  M = 128 rows per CTA (rows of the flattened [clip][103] sequence; rows 101, 102 of a clip give
// garbage frames that are not stored), K = 3 x 160 samples = 15 MMAs of K = 32, N = 128 = 60 bins x (Re, Im) interleaved
// (two N halves: bins 1..60 and 61..119), so that one epilogue thread holds Re and Im of its bins together.
//
// Pairs.  Two CTAs form a cluster and issue cta_group::2 MMAs (M = 256): each CTA stages only its 64 columns of every
// twiddle-digit slab (30 KB instead of 60 KB per slab), which halves both the L2 -> shared traffic of the B stream
// (12 slabs per tile) and the shared-memory operand reads per MMA - a single CTA at N = 128 would run exactly at the
// 128 B/clk shared-memory limit before any fill traffic.
//
// Roles per CTA (12 warps): warp 0 bulk-copy producer of the B slabs, warp 1 MMA issuer (leader CTA; in the peer CTA it
// relays "slab landed" to the leader's barrier), warps 2-3 build the A tile (int16 -> hi / lo byte planes), warps 4-11
// epilogue: TMEM -> 64-bit integer Horner state in registers (64 columns per thread) -> float32 complex64 rounding -> |.|^2
// in float32 -> power spectrum P[frame][120] in global memory.  fe3_post_kernel then does mel -> log -> DCT.  The epilogue
// uses no FP64 instruction: next to a busy tensor pipe they cost 14x their normal issue time (profiles/r04_pipe_probe.txt).
#include "frontend.cuh"
#include "umma.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

namespace honk {

using namespace umma;

constexpr int F3_THREADS = 384;             // warpgroup 0: producer, MMA / relay, two builder warps; warpgroups 1-2: epilogue
constexpr int F3_NBUILD = 64;               // threads that build the A tile (warps 2-3)
constexpr int F3_EPT = (160 + F3_NBUILD - 1) / F3_NBUILD;   // edge samples per builder thread and row
constexpr int F3_J = 6;                    // twiddle digits (F = 6 + 8 (J - 1) = 46 fractional bits)
constexpr int F3_F = 6 + 8 * (F3_J - 1);
constexpr int F3_NSTAGE = 4;               // B slabs in flight
constexpr int F3_NACC = 3;                 // accumulator buffers of 128 TMEM columns; the last 128 columns hold Horner state
constexpr int F3_NF = 480, F3_HOP = 160, F3_HALF = 240;
constexpr int F3_NCHUNK = F3_NF / 16;      // 30 K chunks of 16 samples per frame
constexpr int F3_SEGCH = F3_HOP / 16;      // 10 chunks per segment row
constexpr int F3_AROWS = 130;              // 128 frames + 2 following segment rows
constexpr int F3_APLANE = F3_AROWS * 16;   // one K chunk of all rows
constexpr int F3_ALIMB = F3_SEGCH * F3_APLANE;
constexpr int F3_ABUF = 2 * F3_ALIMB;      // hi limb planes, then lo limb planes
constexpr int F3_NCOLH = 64;               // B columns per CTA per half
constexpr int F3_SLAB = F3_NCHUNK * F3_NCOLH * 16;   // 30720 B: one digit of one N half for one CTA
constexpr int F3_BINS = 119, F3_PSTRIDE = 120;       // power spectrum row: bins 1..119 (+ 1 pad float)
constexpr int F3_SMEM = 2 * F3_ABUF + F3_NSTAGE * F3_SLAB + 1024;

struct Fe3Args {
  const short* pcm;
  long long clip_stride;      // samples
  int n_samples, n_frames, rpc;   // rpc = n_frames + 2 segment rows per clip
  long long total_rows;       // n_clips * rpc
  int n_tilepairs;
  const unsigned char* bslab; // [2 halves][J][2 ranks][F3_SLAB]
  float* P;                   // [n_clips * n_frames][F3_PSTRIDE]
};

namespace {

#ifdef F3_PROFILE
#define F3_T(var) const long long var = clock64()
#define F3_ACC(dst, t0, t1) dst += (t1) - (t0)
#else
#define F3_T(var)
#define F3_ACC(dst, t0, t1)
#endif

__device__ __forceinline__ uint32_t f3_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void f3_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__host__ __device__ constexpr uint32_t f3_idesc(int M, int N, int a_signed) {
  return (2u << 4) | (static_cast<uint32_t>(a_signed) << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) |
         (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void f3_mma2(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives (once all MMAs issued so far by this thread are complete) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void f3_commit2(uint32_t mbar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(mbar),
               "h"(static_cast<uint16_t>(3))
               : "memory");
}
__device__ __forceinline__ void f3_tmem_alloc2(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void f3_tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void f3_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(mbar)
               : "memory");
}
// arrive on the barrier at the same offset in CTA `cta` of the cluster (release at cluster scope)
__device__ __forceinline__ void f3_arrive_cta(uint32_t local_bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(local_bar),
      "r"(cta)
      : "memory");
}
// the same signal without release semantics at cluster scope: for barriers that publish no memory (a TMEM accumulator that
// has been read into registers, a bulk copy that has landed through the async proxy).  A release.cluster arrive costs
// ~900 clk here (it waits for the warp's outstanding global stores), this one a few.
__device__ __forceinline__ void f3_arrive_cta_relaxed(uint32_t local_bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(local_bar),
      "r"(cta)
      : "memory");
}
// wait with acquire at cluster scope (the arrivals may come from the peer CTA); bounded like every wait in this library
__device__ __forceinline__ void f3_wait_cluster(uint32_t mbar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t i = 0; !ok; ++i) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(mbar), "r"(parity)
        : "memory");
    if (i > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ void f3_tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
               "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
               "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
               : "memory");
}
__device__ __forceinline__ void f3_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t f3_elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
__device__ __forceinline__ void f3_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// float32(T 2^-61) for T = 256 S + a, rounded to nearest even ONCE from the exact integer - the complex64 store of librosa's
// stft applied to the exact spectrum - in INTEGER arithmetic plus one int32 -> float32 conversion.  No FP64: while the
// tensor pipe runs, a DFMA / DADD / FP64 conversion of another warp costs ~57 clk instead of ~4 (scripts/probes/pipe_probe.cu,
// profiles/r04_pipe_probe.txt) and slows the MMAs by 11 %.
//   T = 256 V + r,  V = S + floor(a / 256),  0 <= r < 256.
// The 64-bit value Z that is normalised is T itself when V fits an int32 (|T| < 2^39), else V with r as a sticky bit (scale
// 2^8 more).  Z is shifted right by k bits until it fits an int32 (k = 0 when it already does: the conversion is then exact
// up to its own rounding); bits shifted out only matter as a sticky bit OR-ed into bit 0, which then lies at least 7 bits
// below the rounding position of the conversion (also right for negative Z in two's complement: the value moves towards
// +inf by less than one unit, which is what the lost bits stand for).
__device__ __forceinline__ float f3_round_f32(long long S, uint32_t a) {
  const long long V = S + static_cast<long long>(static_cast<int>(a) >> 8);
  const uint32_t r = a & 255u;
  const uint32_t vlo = static_cast<uint32_t>(V);
  const int32_t vhi = static_cast<int32_t>(V >> 32);
  const bool fits = vhi == (static_cast<int32_t>(vlo) >> 31);
  const int32_t hi = fits ? (static_cast<int32_t>(vlo) >> 24) : vhi;
  const uint32_t lo = fits ? ((vlo << 8) | r) : vlo;
  const uint32_t st = fits ? 0u : r;
  const int cls = __clz(hi ^ (hi >> 31));                                   // leading bits of hi equal to its sign: 1..32
  const int k = cls < 32 ? 33 - cls : static_cast<int>((lo >> 31) ^ (static_cast<uint32_t>(hi) & 1u));
  uint32_t lost;                                                             // bits of lo shifted out (shl clamps: k = 0 -> 0)
  asm("shl.b32 %0, %1, %2;" : "=r"(lost) : "r"(lo), "r"(32 - k));
  const int32_t W = static_cast<int32_t>(__funnelshift_r(lo, static_cast<uint32_t>(hi), k)) | static_cast<int32_t>((lost | st) != 0u);
  const int e = k + (fits ? -61 : -53);
  return __int2float_rn(W) * __int_as_float((127 + e) << 23);
}

// The same rounding with the 64-bit integer -> float32 conversion of the XU pipe (I2F.F32.S64, round to nearest even), which
// runs beside the ALU pipe the Horner steps saturate: Z = T when it fits 64 bits (|V| < 2^55), else V with r as a sticky
// bit (V then has at least 56 significant bits, so bit 0 lies far below the rounding position).  -DF3_ROUND_ALU selects the
// all-ALU version above (measured slower: profiles/r04_fe3_round.txt).
__device__ __forceinline__ float f3_round_f32_xu(long long S, uint32_t a) {
  const long long V = S + static_cast<long long>(static_cast<int>(a) >> 8);
  const uint32_t r = a & 255u;
  const int32_t vhi = static_cast<int32_t>(V >> 32);
  const bool fits = static_cast<uint32_t>(vhi + (1 << 23)) < (1u << 24);                 // |V| < 2^55
  const long long Z = fits ? ((V << 8) | static_cast<long long>(r)) : (V | static_cast<long long>(r != 0u));
  return __ll2float_rn(Z) * __int_as_float((127 + (fits ? -61 : -53)) << 23);
}
#ifdef F3_ROUND_ALU
#define F3_ROUND f3_round_f32
#else
#define F3_ROUND f3_round_f32_xu
#endif

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(F3_THREADS, 1) fe3_dft_kernel(const __grid_constant__ Fe3Args a) {
  extern __shared__ __align__(1024) unsigned char f3_smem[];
  __shared__ __align__(8) uint64_t bar_bfull[F3_NSTAGE], bar_bempty[F3_NSTAGE], bar_afull[2], bar_aempty[2];
  __shared__ __align__(8) uint64_t bar_accfull[F3_NACC], bar_accempty[F3_NACC];
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const uint32_t rank = f3_cluster_rank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  unsigned char* sA = f3_smem;                              // [2 bufs][2 limbs][10 chunks][130 rows][16 B]
  unsigned char* sB = f3_smem + 2 * F3_ABUF;                // [NSTAGE][30 chunks][64 cols][16 B]
  const uint32_t a_bfull = smem_u32(&bar_bfull[0]), a_bempty = smem_u32(&bar_bempty[0]);
  const uint32_t a_afull = smem_u32(&bar_afull[0]), a_aempty = smem_u32(&bar_aempty[0]);
  const uint32_t a_accfull = smem_u32(&bar_accfull[0]), a_accempty = smem_u32(&bar_accempty[0]);

  if (tid == 0) {
    for (int s = 0; s < F3_NSTAGE; ++s) {
      mbar_init(a_bfull + 8 * s, leader ? 2 : 1);           // leader: own copy (expect_tx) + the peer's relay
      mbar_init(a_bempty + 8 * s, 1);                       // one multicast commit
    }
    for (int k = 0; k < 2; ++k) {
      mbar_init(a_afull + 8 * k, 2);                        // (leader only) one arrival per CTA's tile builders
      mbar_init(a_aempty + 8 * k, 1);
    }
    for (int k = 0; k < F3_NACC; ++k) {
      mbar_init(a_accfull + 8 * k, 1);
      mbar_init(a_accempty + 8 * k, 16);                    // (leader only) 8 epilogue warps of each CTA
    }
    fence_mbar_init();
  }
  if (warp == 1) f3_tmem_alloc2(smem_u32(&tmem_base_s), 512);
  fence_before();
  __syncthreads();
  f3_cluster_sync();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  const int my_iters = (a.n_tilepairs - pair + n_pairs - 1) / n_pairs;     // tile pairs of this CTA pair

  if (warp == 0) {
    // ================================ B producer (both CTAs) ================================
    if (lane == 0) {
      uint32_t n = 0;
      [[maybe_unused]] long long pw_w = 0;
      for (int it = 0; it < my_iters; ++it)
        for (int h = 0; h < 2; ++h)
          for (int j = 0; j < F3_J; ++j, ++n) {
            const uint32_t s = n % F3_NSTAGE;
            F3_T(q0);
            if (n >= F3_NSTAGE) mbar_wait(a_bempty + 8 * s, ((n / F3_NSTAGE) & 1) ^ 1);
            F3_T(q1);
            F3_ACC(pw_w, q0, q1);
            mbar_expect_tx(a_bfull + 8 * s, F3_SLAB);
            f3_bulk_load(smem_u32(sB + s * F3_SLAB), a.bslab + (static_cast<size_t>(h * F3_J + j) * 2 + rank) * F3_SLAB, F3_SLAB,
                         a_bfull + 8 * s);
          }
#ifdef F3_PROFILE
      if (blockIdx.x < 2) printf("fe3 producer (cta %d): wait_b_empty %lld\n", blockIdx.x, pw_w);
#endif
    }
  } else if (warp == 1) {
    if (!leader) {
      // ================================ relay (peer CTA) ================================
      // the leader's MMAs read this CTA's slab too: tell its barrier when our copy has landed
      if (lane == 0) {
        const uint32_t total = static_cast<uint32_t>(my_iters) * 2 * F3_J;
        for (uint32_t n = 0; n < total; ++n) {
          const uint32_t s = n % F3_NSTAGE;
          mbar_wait(a_bfull + 8 * s, (n / F3_NSTAGE) & 1);
          f3_arrive_cta_relaxed(a_bfull + 8 * s, 0);
        }
      }
    } else {
      // ================================ MMA issuer (leader CTA) ================================
      // The whole warp runs the loops (uniform control flow keeps the descriptor arithmetic on the uniform datapath instead
      // of the vector ALU, which the epilogue warps of this scheduler saturate); one elected lane issues.
      {
        constexpr uint32_t ID_H = f3_idesc(256, 128, 1), ID_L = f3_idesc(256, 128, 0);
        const uint64_t a_hi = make_desc(0, F3_APLANE, 128) & 0xFFFFFFFFFFFF0000ull;    // LBO / SBO / version, address added below
        const uint64_t b_hi = make_desc(0, F3_NCOLH * 16, 128) & 0xFFFFFFFFFFFF0000ull;
        uint32_t n0 = 0, c = 0;                        // first slab of the current (tile, half); accumulator use counter
        [[maybe_unused]] long long pw_a = 0, pw_acc = 0, pw_b = 0;
        F3_T(q_start);
        for (int it = 0; it < my_iters; ++it) {
          const int buf = it & 1;
          F3_T(q0);
          f3_wait_cluster(a_afull + 8 * buf, (it >> 1) & 1);
          F3_T(q1);
          F3_ACC(pw_a, q0, q1);
          fence_after();
          const uint32_t a_h = smem_u32(sA + buf * F3_ABUF), a_l = a_h + F3_ALIMB;
          for (int h = 0; h < 2; ++h, n0 += F3_J) {
            for (int i = 1; i <= F3_J + 1; ++i, ++c) {
              const uint32_t ab = c % F3_NACC;
              if (c >= F3_NACC) {
                F3_T(q2);
                f3_wait_cluster(a_accempty + 8 * ab, ((c / F3_NACC) & 1) ^ 1);
                F3_T(q3);
                F3_ACC(pw_acc, q2, q3);
                fence_after();
              }
              const uint32_t d = tmem + ab * 128;
              uint32_t acc = 0;
              if (i >= 2) {                            // xl . b_(i-1): the slab is resident since the previous level
                const uint32_t nl = n0 + (i - 2), sl = nl % F3_NSTAGE;
                const uint32_t bb = smem_u32(sB + sl * F3_SLAB);
                if (f3_elect_one()) {
#pragma unroll
                  for (int m = 0; m < 15; ++m) {
                    const int seg = m / 5, ks = m - seg * 5;
                    const uint64_t ad = a_hi | ((a_l + (2 * ks) * F3_APLANE + seg * 16) >> 4);
                    const uint64_t bd = b_hi | ((bb + ((seg * 5 + ks) * 2) * (F3_NCOLH * 16)) >> 4);
                    f3_mma2(d, ad, bd, ID_L, m > 0);
                  }
                  f3_commit2(a_bempty + 8 * sl);       // both CTAs may refill this stage
                }
                __syncwarp();
                acc = 1;
              }
              if (i <= F3_J) {                         // xh . b_i
                const uint32_t nh = n0 + (i - 1), sh = nh % F3_NSTAGE;
                F3_T(q4);
                f3_wait_cluster(a_bfull + 8 * sh, (nh / F3_NSTAGE) & 1);
                F3_T(q5);
                F3_ACC(pw_b, q4, q5);
                fence_after();
                const uint32_t bb = smem_u32(sB + sh * F3_SLAB);
                if (f3_elect_one()) {
#pragma unroll
                  for (int m = 0; m < 15; ++m) {
                    const int seg = m / 5, ks = m - seg * 5;
                    const uint64_t ad = a_hi | ((a_h + (2 * ks) * F3_APLANE + seg * 16) >> 4);
                    const uint64_t bd = b_hi | ((bb + ((seg * 5 + ks) * 2) * (F3_NCOLH * 16)) >> 4);
                    f3_mma2(d, ad, bd, ID_H, acc | (m > 0));
                  }
                }
                __syncwarp();
              }
              if (f3_elect_one()) {
                f3_commit2(a_accfull + 8 * ab);        // level i of this half is complete in both CTAs' TMEM
                if (h == 1 && i == F3_J + 1) f3_commit2(a_aempty + 8 * buf);   // the A tiles of both CTAs may be rebuilt
              }
              __syncwarp();
            }
          }
        }
#ifdef F3_PROFILE
        if (blockIdx.x == 0 && lane == 0) printf("fe3 mma: iters %d total %lld wait_a %lld wait_acc_empty %lld wait_b_full %lld\n", my_iters, clock64() - q_start, pw_a, pw_acc, pw_b);
#endif
      }
    }
  } else if (warp < 4) {
    // ================================ A tile builders (both CTAs, two warps) ================================
    const int t64 = tid - 64;                                         // index within the builder group
    const bool fast = (a.clip_stride % 8 == 0) && ((reinterpret_cast<unsigned long long>(a.pcm) & 15) == 0);
    [[maybe_unused]] long long pw_w = 0, pw_build = 0, pw_p1 = 0, pw_p2 = 0;
    for (int it = 0; it < my_iters; ++it) {
      const int buf = it & 1;
      const long long tile = 2ll * (pair + static_cast<long long>(it) * n_pairs) + rank;
      const long long row0 = tile * 128;
      F3_T(q0);
      if (it >= 2) mbar_wait(a_aempty + 8 * buf, ((it >> 1) & 1) ^ 1);
      F3_T(q1);
      F3_ACC(pw_w, q0, q1);
      unsigned char* dst_h = sA + buf * F3_ABUF;
      unsigned char* dst_l = dst_h + F3_ALIMB;
      // The ALU pipe of this warp's scheduler is shared with two epilogue warps, so the builder counts instructions: a
      // thread keeps one chunk column (ch = t % 10) and one of six row phases (t / 10) for the whole tile - no division per
      // item, (clip, u) advance incrementally - and the few chunks that reach into the reflect padding are redone per
      // sample in a second pass over the (at most three) clip boundaries of the tile.
      const long long clip0 = row0 / a.rpc;                        // the only 64-bit division per tile
      const int u0 = static_cast<int>(row0 - clip0 * a.rpc);
      const int rows_here = static_cast<int>(min(static_cast<long long>(F3_AROWS), a.total_rows - row0));   // rows beyond the batch are zero
      const bool regular = fast && a.rpc >= 6 && a.n_samples % 16 == 0;     // chunks are wholly inside or wholly outside a clip
      if (t64 < 60) {
        const int rr = t64 / F3_SEGCH, ch = t64 - rr * F3_SEGCH;
        int u = u0 + rr;
        long long clip = clip0;
        while (u >= a.rpc) { u -= a.rpc; ++clip; }
        for (int rowb = 0; rowb < 132; rowb += 24) {               // 4 steps of 6 rows: eight 16-byte loads in flight
          uint4 v0[4], v1[4];
          int st[4];                                               // 0: pass 2 writes it, 1: zeros, 2: data
#pragma unroll
          for (int bq = 0; bq < 4; ++bq) {
            const int row = rowb + 6 * bq + rr;
            st[bq] = 0;
            v0[bq] = v1[bq] = make_uint4(0, 0, 0, 0);
            if (row < F3_AROWS) {
              if (row >= rows_here) {
                st[bq] = 1;
              } else {
                const int s0 = -F3_HALF + F3_HOP * u + 16 * ch;     // first of this chunk's 16 samples
                if (regular && s0 >= 0 && s0 + 16 <= a.n_samples) {
                  const short* src = a.pcm + clip * a.clip_stride + s0;
#ifdef F3_NOLOAD
                  v0[bq] = make_uint4(row, 1, 2, 3); v1[bq] = v0[bq];
#else
                  v0[bq] = __ldg(reinterpret_cast<const uint4*>(src));
                  v1[bq] = __ldg(reinterpret_cast<const uint4*>(src + 8));
#endif
                  st[bq] = 2;
                }
              }
            }
            u += 6;
            if (u >= a.rpc) { u -= a.rpc; ++clip; }
          }
#pragma unroll
          for (int bq = 0; bq < 4; ++bq) {
            if (st[bq] == 0) continue;
            const int row = rowb + 6 * bq + rr;
            // bytes of a little-endian int16 pair: (lo0, hi0, lo1, hi1) -> gather the low and the high bytes
            uint4 lo, hi;
            lo.x = __byte_perm(v0[bq].x, v0[bq].y, 0x6420); lo.y = __byte_perm(v0[bq].z, v0[bq].w, 0x6420);
            lo.z = __byte_perm(v1[bq].x, v1[bq].y, 0x6420); lo.w = __byte_perm(v1[bq].z, v1[bq].w, 0x6420);
            hi.x = __byte_perm(v0[bq].x, v0[bq].y, 0x7531); hi.y = __byte_perm(v0[bq].z, v0[bq].w, 0x7531);
            hi.z = __byte_perm(v1[bq].x, v1[bq].y, 0x7531); hi.w = __byte_perm(v1[bq].z, v1[bq].w, 0x7531);
            *reinterpret_cast<uint4*>(dst_h + ch * F3_APLANE + row * 16) = hi;
            *reinterpret_cast<uint4*>(dst_l + ch * F3_APLANE + row * 16) = lo;
          }
        }
      }
      F3_T(qp1);
      F3_ACC(pw_p1, q1, qp1);
      // Pass 2, per sample.  Regular case: around every clip boundary inside the tile the rows u = rpc-2, rpc-1, 0, 1 reach
      // into the padding - 4 rows x 160 samples, of which pass 1 wrote the chunks that lie inside the clip.  Otherwise
      // (unaligned buffer, ragged or very short clips) every row is done here.
      auto sample_pass = [&](int row_lo, int row_hi) {               // rows [row_lo, row_hi) of the tile, all 160 samples each
        const int n_s = (row_hi - row_lo) * F3_HOP;
        for (int i0 = t64; i0 < n_s; i0 += 8 * F3_NBUILD) {
          uint32_t val[8];
#pragma unroll
          for (int m = 0; m < 8; ++m) {
            const int i = i0 + m * F3_NBUILD;
            val[m] = 0xffffffffu;                                    // skip
            if (i >= n_s) continue;
            const int row = row_lo + i / F3_HOP, sidx = i % F3_HOP;
            const int x = u0 + row, dc = x / a.rpc, u = x - dc * a.rpc;
            const int s0c = -F3_HALF + F3_HOP * u + (sidx & ~15);
            if (regular && s0c >= 0 && s0c + 16 <= a.n_samples) continue;                        // written by pass 1
            int q = -F3_HALF + F3_HOP * u + sidx;
            if (q < 0) q = -q;                                        // np.pad(mode="reflect"), no edge repeat
            if (q >= a.n_samples) q = 2 * (a.n_samples - 1) - q;
#ifdef F3_NOLOAD
            val[m] = q & 0xffff;
#else
            val[m] = (q >= 0 && q < a.n_samples) ? static_cast<uint16_t>(__ldg(a.pcm + (clip0 + dc) * a.clip_stride + q)) : 0u;
#endif
          }
#pragma unroll
          for (int m = 0; m < 8; ++m) {
            if (val[m] == 0xffffffffu) continue;
            const int i = i0 + m * F3_NBUILD;
            const int row = row_lo + i / F3_HOP, sidx = i % F3_HOP;
            const int off = (sidx >> 4) * F3_APLANE + row * 16 + (sidx & 15);
            dst_l[off] = static_cast<unsigned char>(val[m] & 0xffu);
            dst_h[off] = static_cast<unsigned char>(val[m] >> 8);
          }
        }
      };
      if (regular) {
        // first row of the tile with u == rpc - 2 (or the tile's start if it begins inside an edge group)
        for (int b = a.rpc - u0; b - 2 < rows_here; b += a.rpc)      // b = tile row of a clip's first row (u = 0); may be <= 1
          sample_pass(max(b - 2, 0), min(b + 2, rows_here));
        if (u0 < 2) sample_pass(0, min(2 - u0, rows_here));          // the tile starts inside the head rows of a clip
      } else {
        sample_pass(0, rows_here);
      }
      F3_T(qp2);
      F3_ACC(pw_p2, qp1, qp2);
      fence_async_smem();                                            // generic-proxy stores -> visible to the MMA's async proxy
      f3_bar_sync(1, F3_NBUILD);
      if (t64 == 0) f3_arrive_cta(a_afull + 8 * buf, 0);
      F3_T(q2);
      F3_ACC(pw_build, q1, q2);
    }
#ifdef F3_PROFILE
    if (blockIdx.x < 2 && t64 == 0) printf("fe3 build (cta %d): wait_a_empty %lld build %lld (pass1 %lld pass2 %lld)\n", blockIdx.x, pw_w, pw_build, pw_p1, pw_p2);
#endif
  } else {
    // ================================ epilogue (both CTAs, 8 warps) ================================
    const int ew = warp - 4, quad = warp & 3, cs = ew >> 2;           // TMEM lane quadrant of this warp (= warp % 4); column half
    const int rowl = quad * 32 + lane;
    const uint32_t tbase = tmem + (static_cast<uint32_t>(quad * 32) << 16) + cs * 64;
    const uint32_t thi = tmem + (static_cast<uint32_t>(quad * 32) << 16) + F3_NACC * 128 + cs * 64;   // high words of the Horner state
    static_assert(F3_F + 15 == 61, "final scale");
    uint32_t c = 0;
    [[maybe_unused]] long long pw_w = 0, pw_fin = 0;
    F3_T(q_start);
    for (int it = 0; it < my_iters; ++it) {
      const long long tile = 2ll * (pair + static_cast<long long>(it) * n_pairs) + rank;
      const long long R = tile * 128 + rowl;
      const long long clip = R / a.rpc;
      const int t = static_cast<int>(R - clip * a.rpc);
      const bool store = R < a.total_rows && t < a.n_frames;
      float* prow = a.P + (clip * a.n_frames + t) * F3_PSTRIDE;
      for (int h = 0; h < 2; ++h) {
        // Horner over the levels: T = sum_i a_i 256^(J+1-i).  Levels 1..J in 64-bit INTEGER arithmetic (|a_1| < 2^22 because
        // the top twiddle digit is at most 64, so |sum_{i<=J} a_i 256^(J-i)| < 2^63): an FP64 instruction costs ~57 clk instead
        // of ~4 while the tensor pipe is busy (profiles/r04_pipe_probe.txt), a 64-bit integer shift-add 8 clk either way.  The
        // last level and the float32 rounding are integer work too (f3_round_f32): this kernel has no FP64 instruction.
        // State per value: 64 bits.  The low words stay in registers (64 per thread), the high words live in the last 128
        // TMEM columns of this thread's lane (tcgen05.ld / st per level, 16 at a time): 128 state registers would leave the
        // compiler nothing to work with under the 168-register limit of a 12-warp CTA.
        uint32_t Sl[64];
#pragma unroll 1
        for (int i = 1; i <= F3_J; ++i, ++c) {
          const uint32_t ab = c % F3_NACC;
          F3_T(q0);
          mbar_wait(a_accfull + 8 * ab, (c / F3_NACC) & 1);
          F3_T(q1);
          F3_ACC(pw_w, q0, q1);
          fence_after();
          // two chunks of 16 columns in flight: the loads of chunk q + 1 are issued before the arithmetic of chunk q
          uint32_t v[2][16], sh[2][16];
          tmem_ld16(tbase + ab * 128, v[0]);
          if (i > 1) tmem_ld16(thi, sh[0]);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            tmem_ld_wait();                                          // chunk q has landed
            if (q < 3) {
              tmem_ld16(tbase + ab * 128 + (q + 1) * 16, v[(q + 1) & 1]);
              if (i > 1) tmem_ld16(thi + (q + 1) * 16, sh[(q + 1) & 1]);
            }
            if (i == 1) {
#pragma unroll
              for (int k = 0; k < 16; ++k) {
                Sl[q * 16 + k] = v[q & 1][k];
                sh[q & 1][k] = static_cast<uint32_t>(static_cast<int32_t>(v[q & 1][k]) >> 31);
              }
            } else {
#pragma unroll
              for (int k = 0; k < 16; ++k) {
                const long long S = ((static_cast<long long>(sh[q & 1][k]) << 32) | Sl[q * 16 + k]) * 256 + static_cast<long long>(static_cast<int>(v[q & 1][k]));
                Sl[q * 16 + k] = static_cast<uint32_t>(S);
                sh[q & 1][k] = static_cast<uint32_t>(S >> 32);
              }
            }
            if (q == 3) {                                            // the whole accumulator has been read: hand it back
              fence_before();
              __syncwarp();
              if (lane == 0) f3_arrive_cta_relaxed(a_accempty + 8 * ab, 0);
            }
            f3_tmem_st16(thi + q * 16, sh[q & 1]);
          }
          f3_tmem_st_wait();
        }
        // level J + 1 (xl . b_J), complex64 like librosa's stft output, then |.|^2 in float32
        F3_T(q2);
        const int nb = h == 0 ? 60 : F3_BINS - 60;                   // bins of this half
        const int b0 = cs * 32;                                      // first bin (within the half) of this thread
        {
          const uint32_t ab = c % F3_NACC;
          mbar_wait(a_accfull + 8 * ab, (c / F3_NACC) & 1);
          fence_after();
          ++c;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t v[16], sh[16];
            tmem_ld16(tbase + ab * 128 + q * 16, v);
            tmem_ld16(thi + q * 16, sh);
            tmem_ld_wait();
            if (q == 3) {
              fence_before();
              __syncwarp();
              if (lane == 0) f3_arrive_cta_relaxed(a_accempty + 8 * ab, 0);
            }
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              float pw[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int kk = g * 8 + 2 * e;
                const float re = F3_ROUND((static_cast<long long>(sh[kk]) << 32) | Sl[q * 16 + kk], v[kk]);
                const float im = F3_ROUND((static_cast<long long>(sh[kk + 1]) << 32) | Sl[q * 16 + kk + 1], v[kk + 1]);
                // np.abs(complex64)**2 is hypotf(re, im)^2 in float32; re*re + im*im with one fused rounding is within 2 float32
                // ulps of it (both are within 1.5 ulp of the exact sum) - 1e-7 relative in a mel energy, 1e-7 absolute in a log
                pw[e] = fmaf(re, re, im * im);
              }
              if (store) {
                const int bin = b0 + q * 8 + g * 4;
                if (bin + 4 <= nb) {
                  *reinterpret_cast<float4*>(prow + 60 * h + bin) = make_float4(pw[0], pw[1], pw[2], pw[3]);
                } else {
#pragma unroll
                  for (int e = 0; e < 4; ++e)
                    if (bin + e < nb) prow[60 * h + bin + e] = pw[e];
                }
              }
            }
          }
        }
        F3_T(q3);
        F3_ACC(pw_fin, q2, q3);
      }
    }
#ifdef F3_PROFILE
    if (blockIdx.x < 2 && lane == 0 && (warp == 4 || warp == 8)) printf("fe3 epilogue (cta %d warp %d): total %lld wait_acc_full %lld finalize %lld\n", blockIdx.x, warp, clock64() - q_start, pw_w, pw_fin);
#endif
  }

  fence_before();
  __syncthreads();
  f3_cluster_sync();
  if (warp == 1) f3_tmem_dealloc2(tmem, 512);
}

// ------------------------------------------------------------------------------------------
// P[frame][120] (float32 power of bins 1..119) -> mel (FP64) -> log where > 0 -> DCT (FP64 DMMA) -> float32 MFCC
// (manage_audio.py:31-42 after the stft).  128 threads work on 16 frames at a time, like the post group of fe2_mfcc_kernel.
// ------------------------------------------------------------------------------------------
constexpr int F3P_FR = 16;

template <int KPC, int NM, int ND>
__global__ void __launch_bounds__(128, 4) fe3_post_kernel(const FeDev t, const float* __restrict__ Pg, long long total_frames, int n_items,
                                                       float* __restrict__ out, unsigned char* __restrict__ x0, int n_frames) {
  extern __shared__ __align__(16) unsigned char f3p_smem[];
  constexpr int LMS = NM + 4;
  double* P = reinterpret_cast<double*>(f3p_smem);                   // [16][KPC] power widened to double
  double* LM = P + F3P_FR * KPC;                                     // [16][LMS]
  double* s_melwT = LM + F3P_FR * LMS;                               // [mel_maxw][NM]
  double2* s_log = reinterpret_cast<double2*>(s_melwT + t.mel_maxw * NM);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < t.mel_maxw * NM; i += 128) s_melwT[i] = __ldg(t.melwT + i);
  for (int i = tid; i < 256; i += 128) reinterpret_cast<double*>(s_log)[i] = __ldg(t.logtab + i);
  constexpr int NG = 128 / NM, FPT = (F3P_FR + NG - 1) / NG;
  constexpr int NKD = NM / 4, NCT = ND / 8, CPW = (NCT + 1) / 2;
  const int m = tid % NM, grp = tid / NM;
  const bool act = grp < NG;
  const int lo = __ldg(t.mello + m), cnt = __ldg(t.melcnt + m);
  const int row = lane >> 2, kq = lane & 3;
  const int ft = warp & 1, ct0 = (warp >> 1) ? CPW : 0, nct = (warp >> 1) ? NCT - CPW : CPW;
  double bd[CPW][NKD];
#pragma unroll
  for (int c = 0; c < CPW; ++c)
#pragma unroll
    for (int ks = 0; ks < NKD; ++ks) bd[c][ks] = c < nct ? __ldg(t.dct + (8 * (ct0 + c) + row) * NM + 4 * ks + kq) : 0.0;
  __syncthreads();
  static_assert((F3P_FR * KPC) % 128 == 0, "whole rounds of the 128 threads");
  constexpr int NLD = F3P_FR * KPC / 128;
  float pv[NLD];                                                     // the NEXT item's power values: loaded under this item's arithmetic
  auto fetch = [&](int item) {
    const long long g0 = static_cast<long long>(item) * F3P_FR;
    const long long lim = (total_frames - g0) * KPC;                 // the item's frames are contiguous rows of Pg
#pragma unroll
    for (int u = 0; u < NLD; ++u) pv[u] = (item < n_items && tid + u * 128 < lim) ? __ldg(Pg + g0 * KPC + tid + u * 128) : 0.f;
  };
  fetch(blockIdx.x);
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const long long g0 = static_cast<long long>(item) * F3P_FR;
#pragma unroll
    for (int u = 0; u < NLD; ++u) P[tid + u * 128] = fe2_f32_to_f64(pv[u]);
    fetch(item + gridDim.x);
    __syncthreads();
    if (act) {
      const double* Pb = P + lo;
      double acc[FPT];
#pragma unroll
      for (int q = 0; q < FPT; ++q) acc[q] = 0.0;
      for (int jj = 0; jj < cnt; ++jj) {
        const double w = s_melwT[jj * NM + m];
#pragma unroll
        for (int q = 0; q < FPT; ++q) {
          const int f = grp + q * NG;
          if (f < F3P_FR) acc[q] = fma(w, Pb[f * KPC + jj], acc[q]);
        }
      }
      double lg[FPT];
#pragma unroll
      for (int q = 0; q < FPT; ++q) lg[q] = fe2_log(acc[q], s_log);
#pragma unroll
      for (int q = 0; q < FPT; ++q) {
        const int f = grp + q * NG;
        if (acc[q] > 0.0 && fe2_log_special(acc[q])) lg[q] = log(acc[q]);
        if (f < F3P_FR) LM[f * LMS + m] = acc[q] > 0.0 ? lg[q] : acc[q];   // data[data > 0] = np.log(...) (manage_audio.py:39)
      }
    }
    __syncthreads();
    {
      double o[CPW][2];
#pragma unroll
      for (int c = 0; c < CPW; ++c) o[c][0] = o[c][1] = 0.0;
      const double* La = LM + (8 * ft + row) * LMS + kq;
#pragma unroll
      for (int ks = 0; ks < NKD; ++ks) {
        const double av = La[4 * ks];
#pragma unroll
        for (int c = 0; c < CPW; ++c)
          if (c < nct) fe2_dmma(o[c][0], o[c][1], av, bd[c][ks]);
      }
      const long long g = g0 + 8 * ft + row;
      if (g < total_frames) {
        if (x0 == nullptr) {
          float* op = out + g * ND + 2 * kq;
#pragma unroll
          for (int c = 0; c < CPW; ++c)
            if (c < nct) *reinterpret_cast<float2*>(op + 8 * (ct0 + c)) = make_float2(static_cast<float>(o[c][0]), static_cast<float>(o[c][1]));
        } else {
          // conv0's A operand of the persistent res kernel (res_tc2.cu): x0[clip][w][H + 2][8 halves] = (hi, lo, hi, 0...) of the
          // float32 MFCC, a zero row above and below (conv0's time padding).  Eight consecutive frames of a warp's rows are
          // 128 contiguous bytes per coefficient.
          const long long clip = g / n_frames;
          const int tt = static_cast<int>(g - clip * n_frames);
          uint4* xb = reinterpret_cast<uint4*>(x0) + clip * ND * (n_frames + 2) + 1 + tt;
#pragma unroll
          for (int c = 0; c < CPW; ++c)
            if (c < nct) {
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int w = 8 * (ct0 + c) + 2 * kq + e;
                const float xv = fminf(fmaxf(static_cast<float>(o[c][e]), -60000.f), 60000.f);      // stays finite in f16
                const __half hi = __float2half_rn(xv);
                const __half lo = __float2half_rn(xv - __half2float(hi));
                const uint32_t uh = __half_as_ushort(hi), ul = __half_as_ushort(lo);
                uint4* dst = xb + static_cast<long long>(w) * (n_frames + 2);
                *dst = make_uint4(uh | (ul << 16), uh, 0u, 0u);
                if (tt == 0) dst[-1] = make_uint4(0, 0, 0, 0);
                if (tt == n_frames - 1) dst[1] = make_uint4(0, 0, 0, 0);
              }
            }
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
// Scratch of the allocation-free entry point (honk_frontend_mfcc_noalloc): the float32 power spectrum of at most
// HONK_NOALLOC_MAX_FRAMES frames (64 one-second clips), 480 B per frame = 3 102 720 B.  Every GPU that loads the library
// holds this array until the process exits.  In HOST code the name below is only a stand-in for the symbol, not a device
// address: fe3_plan_create asks cudaGetSymbolAddress for the real address on the front end's device and keeps it in the plan.
static __device__ __align__(256) float f3_static_P[HONK_NOALLOC_MAX_FRAMES * F3_PSTRIDE];

struct Fe3Plan {
  float* d_static_P = nullptr;    // device address of f3_static_P on this plan's device (null: the lookup failed)
  unsigned char* d_bslab = nullptr;
  int n_sms = 148;
  int max_pairs = 74;          // co-resident CTA pairs of fe3_dft_kernel
  int post_smem = 0;
};

void fe3_plan_free(Fe3Plan* p) {
  if (!p) return;
  if (p->d_bslab) cudaFree(p->d_bslab);
  delete p;
}

// `win`: the periodic Hann window [480] exactly as the FP64 path uses it (float64, symmetric).
int fe3_plan_create(const double* win, const FeDev& t, int device, Fe3Plan** out) {
  *out = nullptr;
  if (t.NF != F3_NF || t.hop != F3_HOP || t.k_lo != 1 || t.KP != F3_PSTRIDE || t.n_mels != 40 || t.n_dct != 40) return HONK_OK;
  if (const char* e = getenv("HONK_FE_I8"))
    if (!atoi(e)) return HONK_OK;
  cudaDeviceProp prop;
  HONK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (static_cast<int>(prop.sharedMemPerBlockOptin) < F3_SMEM) return HONK_OK;
  std::vector<unsigned char> slab(static_cast<size_t>(2) * F3_J * 2 * F3_SLAB, 0);
  for (int h = 0; h < 2; ++h) {
    const int nb = h == 0 ? 60 : F3_BINS - 60;
    for (int col = 0; col < 128; ++col) {
      const int b = col >> 1, im = col & 1;
      if (b >= nb) continue;                                // padding columns stay zero
      const int k = 1 + 60 * h + b;
      const int r = col / F3_NCOLH, cc = col % F3_NCOLH;
      for (int n = 0; n < F3_NF; ++n) {
        const int ph = static_cast<int>((static_cast<long long>(k) * n) % F3_NF);      // exact argument reduction
        const double ang = 2.0 * M_PI * ph / F3_NF;
        const double v = win[n] * (im ? -std::sin(ang) : std::cos(ang));
        long long q = std::llround(std::ldexp(v, F3_F));
        signed char dig[F3_J];
        for (int j = F3_J - 1; j >= 0; --j) {               // balanced base-256 digits, least significant first
          long long dg = ((q % 256) + 256) % 256;
          if (dg >= 128) dg -= 256;
          dig[j] = static_cast<signed char>(dg);
          q = (q - dg) / 256;
        }
        if (q != 0) return set_error(HONK_ERR_INVALID, "fe3: twiddle digit overflow");
        for (int j = 0; j < F3_J; ++j)
          slab[(static_cast<size_t>(h * F3_J + j) * 2 + r) * F3_SLAB + (static_cast<size_t>(n / 16) * F3_NCOLH + cc) * 16 + n % 16] =
              static_cast<unsigned char>(dig[j]);
      }
    }
  }
  std::unique_ptr<Fe3Plan, void (*)(Fe3Plan*)> p(new Fe3Plan(), fe3_plan_free);
  HONK_CUDA(cudaMalloc(&p->d_bslab, slab.size()));
  HONK_CUDA(cudaMemcpy(p->d_bslab, slab.data(), slab.size(), cudaMemcpyHostToDevice));
  p->n_sms = prop.multiProcessorCount;
  HONK_CUDA(cudaFuncSetAttribute(fe3_dft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
  {
    // CTA pairs that can be resident at once (a GPC with an odd number of usable SMs strands one): the persistent grid must
    // not exceed it, or the stragglers would run as a second wave
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (p->n_sms / 2));
    cfg.blockDim = dim3(F3_THREADS);
    cfg.dynamicSmemBytes = F3_SMEM;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n_clusters = 0;
    HONK_CUDA(cudaOccupancyMaxActiveClusters(&n_clusters, fe3_dft_kernel, &cfg));
    if (n_clusters < 1) return HONK_OK;
    p->max_pairs = std::min(n_clusters, p->n_sms / 2);
    if (const char* e = getenv("HONK_FE3_PAIRS")) p->max_pairs = std::max(1, std::min(p->max_pairs, atoi(e)));
#ifdef F3_PROFILE
    fprintf(stderr, "fe3: %d co-resident CTA pairs (occupancy query %d)\n", p->max_pairs, n_clusters);
#endif
  }
  p->post_smem = (F3P_FR * F3_PSTRIDE + F3P_FR * 44 + t.mel_maxw * 40) * 8 + 128 * 16;
  HONK_CUDA(cudaFuncSetAttribute(fe3_post_kernel<F3_PSTRIDE, 40, 40>, cudaFuncAttributeMaxDynamicSharedMemorySize, p->post_smem));
  {
    // set-up, with this plan's device current (honk_frontend_create selected it): the device address of the static scratch.
    // A failure here must not take the ordinary paths down: only the no-alloc entry point reports it.
    void* sp = nullptr;
    if (cudaGetSymbolAddress(&sp, f3_static_P) == cudaSuccess) p->d_static_P = static_cast<float*>(sp);
    else (void)cudaGetLastError();
  }
  *out = p.release();
  return HONK_OK;
}

int64_t fe3_workspace_bytes(int64_t n_clips, int64_t n_frames) {
  const int64_t chunk = std::min<int64_t>(n_clips, 4096);
  return align_up(chunk * n_frames * F3_PSTRIDE * 4, 256);
}

// int16 PCM -> float32 MFCC through the power-spectrum scratch `ws` (fe3_workspace_bytes)
int fe3_mfcc(const Fe3Plan* p, const FeDev& t, const short* d_pcm, int64_t n_clips, int64_t n_samples, int64_t clip_stride, float* d_mfcc,
             unsigned char* d_x0, void* ws, cudaStream_t s) {
  const int n_frames = static_cast<int>(1 + n_samples / F3_HOP);
  for (int64_t c0 = 0; c0 < n_clips; c0 += 4096) {
    const int64_t nc = std::min<int64_t>(4096, n_clips - c0);
    Fe3Args a;
    a.pcm = d_pcm + c0 * clip_stride;
    a.clip_stride = clip_stride;
    a.n_samples = static_cast<int>(n_samples);
    a.n_frames = n_frames;
    a.rpc = n_frames + 2;
    a.total_rows = nc * a.rpc;
    const long long n_tiles = (a.total_rows + 127) / 128;
    a.n_tilepairs = static_cast<int>((n_tiles + 1) / 2);
    a.bslab = p->d_bslab;
    a.P = static_cast<float*>(ws);
    {
      ProfSpan span("fe_dft", s);
      const int pairs = std::min(a.n_tilepairs, p->max_pairs);
      fe3_dft_kernel<<<2 * pairs, F3_THREADS, F3_SMEM, s>>>(a);
      HONK_LAUNCH_CHECK();
    }
    {
      ProfSpan span("fe_post", s);
      const long long total_frames = nc * n_frames;
      const int n_items = static_cast<int>((total_frames + F3P_FR - 1) / F3P_FR);
      fe3_post_kernel<F3_PSTRIDE, 40, 40><<<std::min(n_items, p->n_sms * 8), 128, p->post_smem, s>>>(
          t, static_cast<const float*>(ws), total_frames, n_items, d_mfcc ? d_mfcc + c0 * n_frames * 40 : nullptr,
          d_x0 ? d_x0 + c0 * 40 * (n_frames + 2) * 16 : nullptr, n_frames);
      HONK_LAUNCH_CHECK();
    }
  }
  return HONK_OK;
}

// The same two kernels as fe3_mfcc with the power spectrum staged in the file-scope static array: no cudaMalloc / cudaFree,
// no event or stream creation, no caller workspace.  Every call on a device shares that array, so calls that overlap on
// the GPU (two streams, two threads) race; nothing in it has to start at zero (every float the post kernel uses is written
// by the DFT kernel of the same call; the 120th pad float of a row is read but multiplied by no mel weight).
int fe3_mfcc_noalloc(const Fe3Plan* p, const FeDev& t, const short* d_pcm, int64_t n_clips, int64_t n_samples, int64_t clip_stride,
                     float* d_mfcc, cudaStream_t s) {
  HONK_CHECK(p->d_static_P, HONK_ERR_CUDA, "honk_frontend_mfcc_noalloc: cudaGetSymbolAddress of the static scratch failed at set-up");
  const int n_frames = static_cast<int>(1 + n_samples / F3_HOP);
  const long long total_frames = n_clips * n_frames;
  HONK_CHECK(total_frames <= HONK_NOALLOC_MAX_FRAMES, HONK_ERR_INVALID,
             "honk_frontend_mfcc_noalloc: %lld clips x %d frames = %lld frames exceed the static scratch (HONK_NOALLOC_MAX_FRAMES = %d)",
             static_cast<long long>(n_clips), n_frames, total_frames, HONK_NOALLOC_MAX_FRAMES);
  Fe3Args a;
  a.pcm = d_pcm;
  a.clip_stride = clip_stride;
  a.n_samples = static_cast<int>(n_samples);
  a.n_frames = n_frames;
  a.rpc = n_frames + 2;
  a.total_rows = n_clips * a.rpc;
  const long long n_tiles = (a.total_rows + 127) / 128;
  a.n_tilepairs = static_cast<int>((n_tiles + 1) / 2);
  a.bslab = p->d_bslab;
  a.P = p->d_static_P;
  fe3_dft_kernel<<<2 * std::min(a.n_tilepairs, p->max_pairs), F3_THREADS, F3_SMEM, s>>>(a);
  HONK_LAUNCH_CHECK();
  const int n_items = static_cast<int>((total_frames + F3P_FR - 1) / F3P_FR);
  fe3_post_kernel<F3_PSTRIDE, 40, 40><<<std::min(n_items, p->n_sms * 8), 128, p->post_smem, s>>>(t, p->d_static_P, total_frames, n_items,
                                                                                               d_mfcc, nullptr, n_frames);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

}  // namespace honk
