// Persistent multi-layer tcgen05 kernel for the un-pooled res* models (res15, res15-narrow):
// conv0 and every dilated 3x3 C->C layer of SpeechResModel.forward (utils/model.py:104-121) in
// ONE launch, with the activations of the clips in flight kept L2-resident.
//
// GEMM formulation ("ring accumulate").  One task = one INPUT column (MFCC-coefficient index w)
// of one clip: M = 128 time positions, K = all input channels, and the three coefficient taps kw
// are stacked into N = 3*NB (144 for 45->48 channels), so one tcgen05.mma is MMA-bound instead of
// shared-memory-operand bound (N = 48 reaches only 54 % of the pipe, scripts/umma_probe.cu).
// Input column c contributes its kw block to OUTPUT column c - (kw-1)*d.  Columns of one residue
// class mod d form a chain; along a chain the three N blocks of input j land on outputs j-1, j,
// j+1, which sit in ADJACENT 48-column blocks of a TMEM ring, so the 3-way shifted add happens
// inside the tensor core accumulator (no cross-lane traffic, no extra TMEM reads).  Every MMA
// accumulates: the epilogue hands a ring slot back zeroed (tcgen05.st), so no task needs a
// separate overwrite MMA for its first-touched block.  Time taps kh are start-address offsets of
// the A operand (rows of the staged column), time dilation = kh*d rows, zero padding = physically
// zero rows between the channel planes in memory.
//
// conv0 (1 -> C channels, model.py:107-111) runs on the tensor cores too, as layer 0 with K = 16:
// the MFCC value x is split into f16 hi + lo and fed as input channels (hi, lo, hi) against weight
// rows (w_hi, w_hi, w_lo), i.e. x*w to ~2^-21 relative with FP32 accumulation.  A small pre-pass
// writes that 16-byte-per-position image (tc2_pack_x0_kernel); the second 8-channel K half of the
// MMA reads whatever finite f16 data sits one plane further in the stage and meets zero weights.
//
// BatchNorm (affine=False, after the residual add, model.py:112-118) is folded exactly.  The stored
// activation is the raw x_i = relu(conv_i) [+ x_{i-2}] (so the residual stream needs no extra
// copy: even layers update their residual buffer in place and two rotating buffers P, Q per clip
// suffice).  bn_i's scale 1/sqrt(var_i+eps) is folded into conv_{i+1}'s weights, and its mean shift
// into two spare input channels (NB-2, NB-1) that hold 1.0 at real positions and 0 in the padding:
// their per-tap weights are the hi/lo f16 halves of -sum_ci W*scale*mean, so exactly the in-range
// taps subtract the mean - the same result as zero padding applied after bn_i.
//
// Scheduling.  A team of 4 CTAs (consecutive blockIdx) owns K = 2 clips at a time and runs them
// through all layers in lockstep (A.L, B.L, A.L+1, ...), so the team-wide "layer L of clip A
// done" flag (a release counter in global memory) is hidden behind the other clips' work.  Each
// CTA owns whole chains (or chain segments with one halo input column per cut).  Roles: warp 0 =
// bulk-copy (TMA) producer, warp 1 = MMA issuer, then NG groups of 4 epilogue warps (TMEM -> ReLU + residual ->
// f16 -> L2; every group drains every output column, 16 accumulator columns = 2 channel groups each, so a
// slot of the 8-slot logical TMEM ring goes back in 1/NG of a whole-column pass), last warp = flag signaller.
// Working set: 37 teams x 2 clips x 2 buffers x 449 KB = 66 MB, L2-resident.
// Pooled models (res8, res26) run on the same kernel from layer 1 with S clips stacked along the time axis
// per 128-row tile (see res_tc2_plan_create).
#include "res_tc.cuh"
#include "umma.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "model.cuh"

namespace honk {

using namespace umma;

constexpr int T2_G = 4;           // CTAs per team
constexpr int T2_NSTAGE = 8;      // A-operand stages (one input column each)
// warp 0 producer, 1 MMA, then NG groups of 4 epilogue warps (NG = 3 for 6 channel groups, 2 for 4: 16 accumulator columns
// per group either way), then the flag signaller
constexpr int t2_groups(int ncg) { return ncg % 3 == 0 ? 3 : 2; }
constexpr int t2_threads(int ncg) { return (3 + 4 * t2_groups(ncg)) * 32; }
constexpr int T2_PAD = 16;        // zero rows above every channel plane (>= max dilation)
constexpr int T2_NTB = 32;        // "task done" barriers (ring): ONE tcgen05.commit per task serves the producer and the epilogue
#ifndef T2_EARLY_RELEASE
#define T2_EARLY_RELEASE 1         // epilogue hands a TMEM slot back before (1) or after (0) the column's arithmetic and stores
#endif
constexpr int T2_MAXSEG = 8;
constexpr int T2_MAXL = 24;       // C->C layers (res26)
constexpr int T2_NDBG = 16;       // debug counters per CTA

struct T2Seg {                    // a run of output columns of one chain, owned by one CTA
  short r;                        // residue class (first column of the chain)
  short j0, n_out;                // chain indices of the outputs
  short jin0, n_in;               // chain indices of the inputs (outputs +- 1, clamped to the chain)
  short pad[3];
};

struct T2Layer {
  int d;                          // dilation
  int flags;                      // bit0: add residual, bit1: last layer (pool, no store), bit2: conv0 (input = x0 image),
                                  // bit3: input = y0 (pooled conv0 output of a stacked group), bit4: residual = y0
  int w_off;                      // byte offset of this layer's packed weights
  int w_bytes;                    // 3 * nkc B tiles
  int n_seg[T2_G];
  T2Seg seg[T2_G][T2_MAXSEG];
};

struct T2Params {
  const unsigned char* wpack;
  const unsigned char* x0;        // [n_clips][W][H+2][8] f16: (hi, lo, hi, 0...) of the MFCCs, zero first/last row
  const unsigned char* y0;        // pooled models: [n_clips] buffers in the activation layout holding avgpool(relu(conv0))
  unsigned char* P;               // [n_slots] activation buffers
  unsigned char* Q;               // [n_slots]
  float* pool;                    // [n_clips][NB]
  unsigned int* flags;            // [n_clips] team-wide layer counters
  long long slot_stride;          // bytes per clip buffer
  long long x0_clip_stride;
  int n_layers, n_clips, n_slots; // n_layers counts the C->C layers; layers[0] is conv0
  int K;                          // clips per team in flight
  int G;                          // CTAs per team (<= T2_G)
  int L0;                         // first layer the kernel runs: 0 = conv0 (un-pooled), 1 = conv1 (pooled: conv0 + pool ran before)
  int S, Hp1;                     // pooled models stack S clips along the time axis with one zero row between them: a "clip"
                                  // of this kernel is such a group, rows h with h % Hp1 == Hp1 - 1 are the separators (never
                                  // stored, so they stay zero = the time padding of both neighbours); un-pooled: S = 1, Hp1 = H + 1
  int H, W;
  int plane_rows;                 // H + T2_PAD
  int col_bytes;                  // NCG * plane_rows * 16
  int stage_stride, wbytes;       // wbytes = size of one weight buffer (largest layer)
  long long* dbg;                 // optional [gridDim][T2_NDBG] cycle counters (HONK_TC2_DEBUG)
  int l2hint;                     // 1: evict_first on the x0 stream, evict_last on activation stores and loads
  int mode;                       // debug kernel only (results are garbage): 1 = MMA issue alone (no copies, waits or epilogue),
                                  // 2 = epilogue without its global loads/stores and math, 3 = epilogue without its TMEM loads/stores
  // Per-layer tables live in the kernel parameter (constant) bank so that the MMA warp's descriptor
  // arithmetic stays on the uniform datapath (no R2UR per tcgen05.mma operand).
  T2Layer layers[T2_MAXL + 1];
};

namespace {

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred;
}
// Hides how a value was computed: the compiler keeps it in a register instead of rematerialising
// the shared-window address arithmetic (S2UR CgaCtaId + shifts) at every barrier operation.
__device__ __forceinline__ uint32_t keep(uint32_t x) {
  asm volatile("mov.b32 %0, %0;" : "+r"(x));
  return x;
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
// L2 eviction-priority hints (HONK_TC2_L2HINT, default on): the activation ring (66 MB) is rewritten in place every two layers
// and should stay resident, while x0 streams through once.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_load_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void stcg16_hint(void* p, const uint4& v, uint64_t pol) {
  asm volatile("st.global.cg.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_relaxed_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Polls with relaxed loads (an acquire load invalidates the SM's L1 on every poll) and acquires once.
__device__ __forceinline__ void wait_flag(const unsigned int* p, unsigned int target) {
  for (uint32_t i = 0; ld_relaxed_gpu(p) < target; ++i) {
    if (i > (1u << 25)) __trap();
  }
  asm volatile("fence.acq_rel.gpu;" ::: "memory");
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
  uint32_t r;    // one instruction; out-of-range values saturate to +-65504 instead of becoming inf
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// relu(y) + r on packed halves (r = 0 when the layer has no residual)
__device__ __forceinline__ uint32_t relu_add_h2(uint32_t y, uint32_t r) {
  uint32_t o;
  asm("{\n\t.reg .b32 t;\n\tmax.f16x2 t, %1, %3;\n\tadd.rn.f16x2 %0, t, %2;\n\t}" : "=r"(o) : "r"(y), "r"(r), "r"(0u));
  return o;
}
// 32 lanes x 16 columns of zeros (hands an accumulator slot back cleared)
__device__ __forceinline__ void tmem_zero16(uint32_t taddr) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tmem_zero8(uint32_t taddr) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(0u) : "memory");
}
// HB = 16 or 24 accumulator columns of this thread's TMEM lane
template <int HB>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&v)[HB]) {
  static_assert(HB == 16 || HB == 24, "half a slot is 16 or 24 columns");
  tmem_ld16(taddr, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
  if (HB == 24) tmem_ld8(taddr + 16, *reinterpret_cast<uint32_t(*)[8]>(&v[16]));
}
template <int HB>
__device__ __forceinline__ void tmem_zero_cols(uint32_t taddr) {
  tmem_zero16(taddr);
  if (HB == 24) tmem_zero8(taddr + 16);
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

#define T2_TIMED(acc, stmt) do { if (DBG) { const long long t__ = clock64(); stmt; acc += clock64() - t__; } else { stmt; } } while (0)

// MMAs [I0, I1) of one task: index i = kh * NK + kc reads the A rows shifted by kh*d and the 16-channel chunk kc, against B
// tile i.  (A task's accumulator slots never straddle the end of TMEM: see NPHYS / NSLOT in the kernel.)
template <int I0, int I1, int NK, int NB>
__device__ __forceinline__ void issue_mmas(uint64_t a_hi, uint32_t a_lo, uint32_t pr2, uint32_t d, uint64_t b_hi, uint32_t b_f,
                                           uint32_t d1, uint32_t id1) {
#pragma unroll
  for (int i = I0; i < I1; ++i) {
    const int kh = i / NK, kc = i - kh * NK;
    const uint64_t ad = a_hi | static_cast<uint64_t>(a_lo + kc * pr2 + kh * d);
    const uint64_t bd = b_hi | static_cast<uint64_t>(b_f + i * 6 * NB);
    mma_f16_ss(d1, ad, bd, id1, 1);
  }
}

// Second half of a task in ONE instruction sequence: the two look-ahead probes are issued first, then (one elected
// lane) four MMAs and up to three commits, and only then are the probe results combined - so the probes' latency
// overlaps the MMA issue instead of stalling this warp between the two halves of a task.
__device__ __forceinline__ uint32_t probe_issue4_commit(uint32_t pa, uint32_t pap, uint32_t pb, uint32_t pbp, uint32_t d1,
                                                        uint32_t id1, uint64_t a0, uint64_t a1, uint64_t a2, uint64_t a3, uint64_t b0,
                                                        uint64_t b1, uint64_t b2, uint64_t b3, uint32_t c_empty, uint32_t c_f1,
                                                        uint32_t c_f2) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p, q, pe, pt, pc1, pc2;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 q, [%3], %4;\n"
      "elect.sync _|pe, 0xFFFFFFFF;\n"
      "setp.eq.u32 pt, %5, %5;\n"
      "setp.ne.and.b32 pc1, %16, 0, pe;\n"
      "setp.ne.and.b32 pc2, %17, 0, pe;\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%5], %7, %11, %6, pt;\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%5], %8, %12, %6, pt;\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%5], %9, %13, %6, pt;\n"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%5], %10, %14, %6, pt;\n"
      "@pe tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%15];\n"
      "@pc1 tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%16];\n"
      "@pc2 tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%17];\n"
      "and.pred p, p, q;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(pa), "r"(pap), "r"(pb), "r"(pbp), "r"(d1), "r"(id1), "l"(a0), "l"(a1), "l"(a2), "l"(a3), "l"(b0), "l"(b1), "l"(b2),
        "l"(b3), "r"(c_empty), "r"(c_f1), "r"(c_f2)
      : "memory");
  return ok;
}

template <int NCG, bool DBG>
__global__ void __launch_bounds__(t2_threads(NCG), 1) tc2_res_kernel(const __grid_constant__ T2Params p) {
  constexpr int NG = t2_groups(NCG), T2_THREADS = t2_threads(NCG), SIG_WARP = 2 + 4 * NG;
  constexpr int NB = NCG * 8;            // padded channels = N block
  constexpr int NKC = NCG / 2;           // K chunks of 16 channels
  constexpr int NPHYS = 512 / NB;        // accumulator slots in TMEM
  constexpr int NSLOT = NPHYS - 2;       // LOGICAL ring of output-column accumulators.  A task writes three consecutive slots
                                         // starting at physical slot (g - 1) % NSLOT <= NSLOT - 1, so it never straddles the end
                                         // of TMEM (no MMA split in two): physical slots NSLOT, NSLOT + 1 are second homes of
                                         // logical slots 0, 1, and the epilogue adds the two homes of those outputs.
  constexpr int NMMA = 3 * NKC;          // MMAs per input column
  constexpr int NSPLIT = NMMA / 2 + 1;   // MMAs issued before the look-ahead probe of the next task

  extern __shared__ __align__(1024) unsigned char smem[];
  // A tcgen05.commit costs the issuing thread ~97 clk and an MMA ~54 clk (scripts/umma2_probe.cu): with the two or three
  // commits per task of a per-stage "empty" and a per-slot "accumulator full" barrier the issue stream (~700 clk) was
  // longer than the 648 clk the 9 MMAs take.  Hence one barrier per TASK, task t -> bar_task[t % T2_NTB]: the producer
  // waits for task it - T2_NSTAGE before refilling its stage, the epilogue for the task that contributes last to a column.
  // A waiter is never a whole ring behind: the MMA warp runs at most T2_NSTAGE tasks ahead of the producer's loads and,
  // through the accumulator ring, fewer than ~10 tasks ahead of the epilogue.
  __shared__ __align__(8) uint64_t bar_full[T2_NSTAGE], bar_task[T2_NTB];
  __shared__ __align__(8) uint64_t bar_acc_empty[NSLOT];
  __shared__ __align__(8) uint64_t bar_wfull[2], bar_wfree[2];
  __shared__ __align__(8) uint64_t bar_done[4];        // epilogue -> signaller: this CTA's stores of one (clip, layer) are issued
  __shared__ __align__(8) uint64_t bar_dbg;            // debug mode 1: pipe drain
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // provably warp-uniform: keeps role code on the uniform datapath
  const uint32_t a_full = keep(smem_u32(&bar_full[0])), a_task = keep(smem_u32(&bar_task[0]));
  const uint32_t a_acc_empty = keep(smem_u32(&bar_acc_empty[0]));
  const uint32_t a_wfull = keep(smem_u32(&bar_wfull[0])), a_wfree = keep(smem_u32(&bar_wfree[0]));
  const uint32_t a_done = keep(smem_u32(&bar_done[0]));
  const int G = p.G, L0 = p.L0;
  const int team = blockIdx.x / G, rank = blockIdx.x % G, n_teams = gridDim.x / G;
  unsigned char* stages = smem;
  unsigned char* wbuf = smem + T2_NSTAGE * p.stage_stride;

  if (warp == 1) tmem_alloc(smem_u32(&tmem_base_s), 512);
  if (tid == 0) {
    for (int s = 0; s < T2_NSTAGE; ++s) mbar_init(a_full + 8 * s, 1);
    for (int k = 0; k < T2_NTB; ++k) mbar_init(a_task + 8 * k, 1);
    for (int k = 0; k < NSLOT; ++k) mbar_init(a_acc_empty + 8 * k, 4 * NG);
    for (int k = 0; k < 2; ++k) {
      mbar_init(a_wfull + 8 * k, 1);
      mbar_init(a_wfree + 8 * k, 1);
    }
    for (int k = 0; k < 4; ++k) mbar_init(a_done + 8 * k, 4 * NG);
    mbar_init(smem_u32(&bar_dbg), 1);
    fence_mbar_init();
  }
  // conv0's second K half reads stage bytes no copy of that layer wrote: they must be finite f16 from the start
  for (int i = tid * 16; i < T2_NSTAGE * p.stage_stride; i += T2_THREADS * 16) *reinterpret_cast<uint4*>(stages + i) = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem = tmem_base_s;
  if (warp >= 2 && warp < 6) {                        // accumulators start from zero: every MMA accumulates
    for (int c = 0; c < 512 / 16; ++c) tmem_zero16(tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + c * 16);
    tmem_st_wait();
  }
  fence_before();
  __syncthreads();
  fence_after();
  const int K = p.K;                                  // clips a team runs in lockstep (layer-interleaved)
  const int n_pairs = (p.n_clips + K - 1) / K;        // clip groups
  const int plane_bytes = p.plane_rows * 16;
  const int x0_col_bytes = (p.H + 2) * 16;

  if (warp == 0) {
    // ======================================= producer =======================================
    if (lane == 0 && !(DBG && p.mode == 1)) {
      uint32_t it = 0, q = 0;
      int piter = 0;
      long long t_flag = 0, t_empty = 0;
      const long long t_start = clock64();
      for (int pair = team; pair < n_pairs; pair += n_teams, ++piter) {
        const int nab = min(K, p.n_clips - K * pair);
        for (int L = L0; L <= p.n_layers; ++L, ++q) {
          const T2Layer& ly = p.layers[L];
          if (q == 0) {
            mbar_expect_tx(a_wfull, ly.w_bytes);
            bulk_load(smem_u32(wbuf), p.wpack + ly.w_off, ly.w_bytes, a_wfull);
          }
          const int d = ly.d;
          const bool from_x0 = ly.flags & 4;
          // x0 is read once; the activation ring is re-read by the next layer and rewritten.  The pooled models' y0 image is read
          // twice (layer 1's input, layer 2's residual) shortly after each other: no hint (evict_first cost 0.6 GB of re-reads)
          const uint64_t in_pol = from_x0 ? l2_policy_evict_first() : l2_policy_evict_last();
          const bool hint_in = p.l2hint && !(ly.flags & 8);
          const uint32_t bytes = from_x0 ? x0_col_bytes : (NCG - 1) * plane_bytes + (p.H + 2 * d) * 16;
          const long long colb = from_x0 ? x0_col_bytes : p.col_bytes;
          const int nseg = ly.n_seg[rank];
          for (int ab = 0; ab < nab; ++ab) {
            const int clip = K * pair + ab, slot = K * team + ab;
            const unsigned char* in;
            if (from_x0) {
              in = p.x0 + static_cast<long long>(clip) * p.x0_clip_stride;
              // conv0 writes Q[slot]: the team's last-layer MMAs of the previous clip in this slot still read it
              if (piter > 0) T2_TIMED(t_flag, wait_flag(p.flags + (clip - K * n_teams), G * (p.n_layers + 1)));
            } else if (ly.flags & 8) {
              in = p.y0 + static_cast<long long>(clip) * p.slot_stride + (T2_PAD - d) * 16;
              // this layer writes P[slot]: the team's last layers of the previous group in this slot still read P and Q
              if (piter > 0) T2_TIMED(t_flag, wait_flag(p.flags + (clip - K * n_teams), G * (p.n_layers + 1 - L0)));
            } else {
              in = ((L & 1) ? p.Q : p.P) + static_cast<long long>(slot) * p.slot_stride + (T2_PAD - d) * 16;
              T2_TIMED(t_flag, wait_flag(p.flags + clip, G * (L - L0)));    // every CTA of the team finished layer L-1
              fence_proxy_async_global();
            }
            for (int s = 0; s < nseg; ++s) {
              const T2Seg& sg = ly.seg[rank][s];
              const int r = sg.r, jin0 = sg.jin0, n_in = sg.n_in;
              for (int t = 0; t < n_in; ++t, ++it) {
                const int st = it % T2_NSTAGE;
                if (it >= T2_NSTAGE) {                     // the task that last read this stage is done
                  const uint32_t tp = it - T2_NSTAGE;
                  T2_TIMED(t_empty, mbar_wait(a_task + 8 * (tp % T2_NTB), (tp / T2_NTB) & 1));
                }
                const uint32_t full = a_full + 8 * st;
                mbar_expect_tx(full, bytes);
                if (hint_in) bulk_load_hint(smem_u32(stages + st * p.stage_stride), in + (r + (jin0 + t) * d) * colb, bytes, full, in_pol);
                else bulk_load(smem_u32(stages + st * p.stage_stride), in + (r + (jin0 + t) * d) * colb, bytes, full);
              }
            }
          }
          // prefetch the weights of the next layer in sequence into the other buffer
          const bool more = (L < p.n_layers) || (pair + n_teams < n_pairs);
          if (more) {
            const uint32_t qn = q + 1;
            const T2Layer& nl = p.layers[L < p.n_layers ? L + 1 : L0];
            if (qn >= 2) mbar_wait(a_wfree + 8 * (qn & 1), ((qn - 2) >> 1) & 1);
            mbar_expect_tx(a_wfull + 8 * (qn & 1), nl.w_bytes);
            bulk_load(smem_u32(wbuf + (qn & 1) * p.wbytes), p.wpack + nl.w_off, nl.w_bytes, a_wfull + 8 * (qn & 1));
          }
        }
      }
      if (DBG) { long long* o = p.dbg + blockIdx.x * T2_NDBG; o[0] = clock64() - t_start; o[1] = t_flag; o[2] = t_empty; }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer =====================================
    const uint64_t a_hi = make_desc(0, plane_bytes, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t a_lbo = (static_cast<uint32_t>(p.plane_rows) & 0x3FFF) << 16;
    const uint64_t b_hi = make_desc(0, 3 * NB * 16, 128) & 0xFFFFFFFF00000000ull;
    const uint32_t b_lbo = (static_cast<uint32_t>(3 * NB) & 0x3FFF) << 16;
    const uint32_t stage16 = static_cast<uint32_t>(p.stage_stride) >> 4;
    const uint32_t a_st0 = a_lbo | (smem_u32(stages) >> 4);
    const uint32_t pr2 = 2 * p.plane_rows;
    constexpr uint32_t IDESC0 = make_idesc(128, 0, 0), IDESC_BLK = static_cast<uint32_t>(NB >> 3) << 17;
    uint32_t st = 0, stpar = 0, q = 0, tk = 0;  // tk = running task index (the producer's `it`, the epilogue's `tk0 + ...`)
    uint32_t a_st = a_st0;                      // low descriptor word of stage st
    int gseg = 2 * NSLOT;                       // ring index of the segment's first output (offset keeps halo indices >= 0)
    long long t_acc = 0, t_full = 0, t_w = 0, n_task = 0, n_hit = 0;
    const long long t_start = clock64();
    for (int pair = team; pair < n_pairs; pair += n_teams) {
      const int nab = min(K, p.n_clips - K * pair);
      for (int L = L0; L <= p.n_layers; ++L, ++q) {
        const T2Layer& ly = p.layers[L];
        const uint32_t d = ly.d;
        const bool conv0 = ly.flags & 4;
        const int nseg = ly.n_seg[rank];
        if (!(DBG && p.mode == 1)) T2_TIMED(t_w, mbar_wait(a_wfull + 8 * (q & 1), (q >> 1) & 1));
        const uint32_t b_lo0 = b_lbo | (smem_u32(wbuf + (q & 1) * p.wbytes) >> 4);
        for (int ab = 0; ab < nab; ++ab) {
          for (int s = 0; s < nseg; ++s) {
            const T2Seg& sg = ly.seg[rank][s];
            const int j0 = sg.j0, n_out = sg.n_out, jin0 = sg.jin0, n_in = sg.n_in;
            const int jend = j0 + n_out - 1;
            int gc = gseg + (jin0 - j0);          // ring index of the output column aligned with the input column
            bool pre = false;                     // this task's barriers were already seen complete by the look-ahead
            for (int t = 0; t < n_in; ++t, ++gc) {
              const int jc = jin0 + t;
              const bool lo_in = jc > j0, c_in = (jc >= j0) && (jc <= jend), hi_in = jc < jend;
              const int first_blk = lo_in ? 0 : (c_in ? 1 : 2), last_blk = hi_in ? 2 : (c_in ? 1 : 0);
              const int n = last_blk - first_blk + 1;
              const uint32_t sf = static_cast<uint32_t>(gc - 1) % NSLOT + first_blk;     // physical slot, <= NSLOT + 1
              // first-touched outputs need their ring slot handed back (read and zeroed) by the epilogue:
              // block 2, and at t == 0 block 1 too
              if (!pre && !(DBG && p.mode == 1)) {
                const long long tw0 = DBG ? clock64() : 0;
                if (t == 0 && c_in) {
                  const uint32_t g1 = static_cast<uint32_t>(gc);
                  mbar_wait(a_acc_empty + 8 * (g1 % NSLOT), ((g1 / NSLOT) & 1) ^ 1);
                }
                if (hi_in) {
                  const uint32_t g2 = static_cast<uint32_t>(gc + 1);
                  mbar_wait(a_acc_empty + 8 * (g2 % NSLOT), ((g2 / NSLOT) & 1) ^ 1);
                }
                const long long tw1 = DBG ? clock64() : 0;
                mbar_wait(a_full + 8 * st, stpar);
                if (DBG) { t_acc += tw1 - tw0; t_full += clock64() - tw1; }
              }
              fence_after();
              const uint32_t b_f = b_lo0 + first_blk * NB;
              const uint32_t d1 = tmem + sf * NB;
              const uint32_t id1 = IDESC0 + n * IDESC_BLK;
              if (elect_one()) {
                if (conv0) {
                  issue_mmas<0, 3, 1, NB>(a_hi, a_st, pr2, d, b_hi, b_f, d1, id1);
                } else {
                  issue_mmas<0, NSPLIT, NKC, NB>(a_hi, a_st, pr2, d, b_hi, b_f, d1, id1);
                }
              }
              __syncwarp();
              // Look one task ahead while the queued MMAs drain: probe (never block on) the barriers the next
              // task of this segment needs, so that its issue can start without a serialized poll.
              pre = false;
              if (NMMA - NSPLIT == 4 && !conv0 && !(DBG && p.mode == 1)) {
                // fast path (every regular task of res15): probes, second-half MMAs and commits in one sequence
                const bool has_next = t + 1 < n_in;
                const uint32_t st1 = (st + 1 == T2_NSTAGE) ? 0 : st + 1;
                const uint32_t par1 = (st + 1 == T2_NSTAGE) ? stpar ^ 1 : stpar;
                const uint32_t g2 = static_cast<uint32_t>(gc + 2);
                const bool touch = jc + 1 < jend;                        // next task first-touches block gc + 2
                const uint32_t pa = has_next ? a_full + 8 * st1 : a_full + 8 * st, pap = has_next ? par1 : stpar;
                const uint32_t pb = (has_next && touch) ? a_acc_empty + 8 * (g2 % NSLOT) : pa;
                const uint32_t pbp = (has_next && touch) ? ((g2 / NSLOT) & 1) ^ 1 : pap;
                uint64_t ad[4], bd[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const int ii = NSPLIT + i, kh = ii / NKC, kc = ii - kh * NKC;
                  ad[i] = a_hi | static_cast<uint64_t>(a_st + kc * pr2 + kh * d);
                  bd[i] = b_hi | static_cast<uint64_t>(b_f + ii * 6 * NB);
                }
                const uint32_t ok = probe_issue4_commit(
                    pa, pap, pb, pbp, d1, id1, ad[0], ad[1], ad[2], ad[3], bd[0], bd[1], bd[2], bd[3], a_task + 8 * (tk % T2_NTB),
                    0u, 0u);
                __syncwarp();
                pre = ok && has_next;
                if (DBG && has_next) { n_task += 1; n_hit += pre; }
              } else {
              if (t + 1 < n_in && !(DBG && p.mode == 1)) {
                const uint32_t st1 = (st + 1 == T2_NSTAGE) ? 0 : st + 1;
                const uint32_t par1 = (st + 1 == T2_NSTAGE) ? stpar ^ 1 : stpar;
                bool ok;
                if (jc + 1 < jend) {                                    // next task first-touches block gc + 2
                  const uint32_t g2 = static_cast<uint32_t>(gc + 2);
                  ok = mbar_test_wait2(a_full + 8 * st1, par1, a_acc_empty + 8 * (g2 % NSLOT), ((g2 / NSLOT) & 1) ^ 1);
                } else {
                  ok = mbar_test_wait(a_full + 8 * st1, par1);
                }
                pre = ok;
                if (DBG) { n_task += 1; n_hit += ok; }
              }
              if (elect_one()) {
                if (!conv0) {
                  issue_mmas<NSPLIT, NMMA, NKC, NB>(a_hi, a_st, pr2, d, b_hi, b_f, d1, id1);
                }
                commit(a_task + 8 * (tk % T2_NTB));
              }
              __syncwarp();
              }
              ++tk;
              a_st += stage16;
              if (++st == T2_NSTAGE) { st = 0; stpar ^= 1; a_st = a_st0; }
            }
            gseg += n_out;
          }
        }
        if (elect_one()) commit(a_wfree + 8 * (q & 1));
        __syncwarp();
      }
    }
    if (DBG && p.mode == 1) {                   // drain the tensor pipe before reading the clock
      if (elect_one()) commit(smem_u32(&bar_dbg));
      __syncwarp();
      mbar_wait(smem_u32(&bar_dbg), 0);
    }
    if (DBG && lane == 0) {
      long long* o = p.dbg + blockIdx.x * T2_NDBG;
      o[3] = clock64() - t_start; o[4] = t_acc; o[5] = t_full; o[6] = t_w; o[11] = n_hit; o[12] = n_task;
    }
  } else if (warp < SIG_WARP && !(DBG && p.mode == 1)) {
    // ======================================= epilogue =======================================
    // Every group of 4 warps drains its share of the channels of every output column.  Stored activations
    // are the raw x_L = relu(conv) [+ x_{L-2}]; BatchNorm lives in the next layer's weights (scale) and in
    // its two constant-one input channels (mean shift, exact at the zero-padded borders).
    // All NG groups of 4 warps work on EVERY output column, each on its share of the channels (HC channel groups = HB
    // accumulator columns): a column is drained in 1/NG of the time of a whole-column pass, which is what the 8-slot ring
    // needs.
    constexpr int HC = NCG / NG, HB = HC * 8;
    const uint64_t st_pol = l2_policy_evict_last();
    const int ew = warp - 2, quad = warp & 3, half = ew >> 2;   // half = this warp's group
    const int cg0 = half * HC;
    const int h = quad * 32 + lane;
    const int sub = h / p.Hp1;                              // clip of a stacked group this row belongs to
    const bool valid = h < p.H && (h - sub * p.Hp1) < p.Hp1 - 1;
    const int etid = ew * 32 + lane;
    uint32_t g = 0, ev = 0, tk0 = 0;            // tk0 = tasks before the current segment (same order as the MMA warp)
    long long t_wait = 0;
    [[maybe_unused]] unsigned long long n_sat = 0;   // DBG: accumulator values beyond the f16 range (stored saturated)
    const long long t_start_e = clock64();
    for (int pair = team; pair < n_pairs; pair += n_teams) {
      const int nab = min(K, p.n_clips - K * pair);
      for (int L = L0; L <= p.n_layers; ++L) {
        const T2Layer& ly = p.layers[L];
        const int d = ly.d, flags = ly.flags;
        const bool add_res = flags & 1, last = flags & 2;
        const int nseg = ly.n_seg[rank];
        const long long cstep = static_cast<long long>(d) * p.col_bytes;   // next column of a segment
        for (int ab = 0; ab < nab; ++ab) {
          const int clip = K * pair + ab, slot = K * team + ab;
          unsigned char* Qs = p.Q + static_cast<long long>(slot) * p.slot_stride + (T2_PAD + h) * 16;
          unsigned char* Ps = p.P + static_cast<long long>(slot) * p.slot_stride + (T2_PAD + h) * 16;
          unsigned char* outb = (L & 1) ? Ps : Qs;            // conv0 -> Q; odd: Q -> P; even: P (+Q residual) -> Q in place
          const unsigned char* resb = (flags & 16) ? p.y0 + static_cast<long long>(clip) * p.slot_stride + (T2_PAD + h) * 16 : Qs;
          const bool ld_res = add_res && valid && !(DBG && p.mode == 2);
          float psum[HB];
          if (last) {
#pragma unroll
            for (int k = 0; k < HB; ++k) psum[k] = 0.f;
          }
          for (int s = 0; s < nseg; ++s) {
            const T2Seg& sg = ly.seg[rank][s];
            const int j0 = sg.j0, jlim = sg.j0 + sg.n_out;
            uint32_t gi = g;
            g += sg.n_out;
            const uint32_t tks = tk0;                          // task index of this segment's first input column
            tk0 += sg.n_in;
            const int jin0 = sg.jin0, jin_last = sg.jin0 + sg.n_in - 1;
            long long off = static_cast<long long>(sg.r + j0 * d) * p.col_bytes + static_cast<long long>(cg0) * plane_bytes;
            uint4 rv[HC];                                      // residual, prefetched one column ahead
#pragma unroll
            for (int c = 0; c < HC; ++c) rv[c] = make_uint4(0, 0, 0, 0);
            if (ld_res) {
#pragma unroll
              for (int c = 0; c < HC; ++c) rv[c] = ldcg16(resb + off + c * plane_bytes);
            }
            for (int o = j0; o < jlim; ++o, ++gi, off += cstep) {
              const uint32_t sl = gi % NSLOT;
              uint4 rc[HC];
#pragma unroll
              for (int c = 0; c < HC; ++c) rc[c] = rv[c];
              if (ld_res && o + 1 < jlim) {
#pragma unroll
                for (int c = 0; c < HC; ++c) rv[c] = ldcg16(resb + off + cstep + c * plane_bytes);
              }
              // column o is complete when the task of input column min(o + 1, last input of the segment) is
              const uint32_t tkc = tks + static_cast<uint32_t>(min(o + 1, jin_last) - jin0);
              T2_TIMED(t_wait, mbar_wait(a_task + 8 * (tkc % T2_NTB), (tkc / T2_NTB) & 1));
              fence_after();
              uint32_t v[HB];
              const uint32_t taddr = tmem + (static_cast<uint32_t>(quad * 32) << 16) + sl * NB + half * HB;
              if (DBG && p.mode == 3) {
#pragma unroll
                for (int c = 0; c < HB; ++c) v[c] = 0;
              } else {
                tmem_ld_cols<HB>(taddr, v);
                if (sl < 2) {                                  // logical slots 0, 1 have a second home past the ring's end
                  uint32_t v2[HB];
                  const uint32_t taddr2 = taddr + NSLOT * NB;
                  tmem_ld_cols<HB>(taddr2, v2);
                  tmem_ld_wait();
                  tmem_zero_cols<HB>(taddr2);
#pragma unroll
                  for (int c = 0; c < HB; ++c) v[c] = __float_as_uint(__uint_as_float(v[c]) + __uint_as_float(v2[c]));
                } else {
                  tmem_ld_wait();
                }
                tmem_zero_cols<HB>(taddr);                     // the slot goes back cleared
                // hand the slot back as soon as the zeros have landed: the values are in registers, the MMA warp need not
                // wait for this column's arithmetic and stores
                if (T2_EARLY_RELEASE) {
                  tmem_st_wait();
                  fence_before();
                  __syncwarp();
                  if (lane == 0) mbar_arrive(a_acc_empty + 8 * sl);
                }
              }
              if (DBG && valid) {                                // f16 storage saturates silently (pack_h2): count what would
#pragma unroll
                for (int c = 0; c < HB; ++c) n_sat += fabsf(__uint_as_float(v[c])) > 65504.f;
              }
              if (DBG && p.mode == 2) {
              } else if (!last) {
#pragma unroll
                for (int c = 0; c < HC; ++c) {
                  uint32_t w4[4];
                  const uint32_t* rr = reinterpret_cast<const uint32_t*>(&rc[c]);
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    w4[k] = relu_add_h2(pack_h2(__uint_as_float(v[c * 8 + 2 * k]), __uint_as_float(v[c * 8 + 2 * k + 1])), rr[k]);
                  if (c == HC - 1 && half == NG - 1) w4[3] = 0x3C003C00u;   // the two constant-one channels (last channel group)
                  if (valid) {
                    if (p.l2hint) stcg16_hint(outb + off + c * plane_bytes, make_uint4(w4[0], w4[1], w4[2], w4[3]), st_pol);
                    else stcg16(outb + off + c * plane_bytes, make_uint4(w4[0], w4[1], w4[2], w4[3]));
                  }
                }
              } else if (valid) {
#pragma unroll
                for (int c = 0; c < HC; ++c) {
                  const __half2* rh = reinterpret_cast<const __half2*>(&rc[c]);
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    const float2 f = __half22float2(rh[k]);
                    psum[c * 8 + 2 * k] += fmaxf(__uint_as_float(v[c * 8 + 2 * k]), 0.f) + f.x;
                    psum[c * 8 + 2 * k + 1] += fmaxf(__uint_as_float(v[c * 8 + 2 * k + 1]), 0.f) + f.y;
                  }
                }
              }
              if (DBG && p.mode == 3) {
                __syncwarp();
                if (lane == 0) mbar_arrive(a_acc_empty + 8 * sl);
              } else if (!T2_EARLY_RELEASE) {                  // the zero stores have had the column's arithmetic to land
                tmem_st_wait();
                fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_acc_empty + 8 * sl);
              }
            }
          }
          if (last) {
            // rows of one warp belong to at most two clips of a stacked group (un-pooled: S = 1, one pass)
            const int s_lo = __shfl_sync(0xffffffffu, sub, 0), s_hi = min(__shfl_sync(0xffffffffu, sub, 31), p.S - 1);
            for (int sc = s_lo; sc <= s_hi; ++sc) {
#pragma unroll
              for (int k = 0; k < HB; ++k) {
                float sum = (sub == sc) ? psum[k] : 0.f;          // invalid rows never accumulated
#pragma unroll
                for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                if (lane == 0) atomicAdd(p.pool + (static_cast<long long>(clip) * p.S + sc) * NB + half * HB + k, sum);
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(a_done + 8 * (ev & 3));     // this warp's stores for (clip, L) are issued
          ++ev;
        }
      }
    }
    if (DBG && etid == 0) { long long* o = p.dbg + blockIdx.x * T2_NDBG; o[7] = t_wait; o[8] = clock64() - t_start_e; }
    if (DBG && n_sat) atomicAdd(reinterpret_cast<unsigned long long*>(p.dbg + blockIdx.x * T2_NDBG + 13), n_sat);
  } else if (warp == SIG_WARP && !(DBG && p.mode == 1)) {
    // ======================================= flag signaller =================================
    // Publishes "all stores of this CTA for (clip, layer) are visible" to the team: waits until the eight
    // epilogue warps have arrived, then a gpu-scope fence (cumulative over the stores observed through the
    // mbarrier) + release increment of the clip's layer counter.  Keeps fences off the epilogue warps.
    if (lane == 0) {
      uint32_t ev = 0;
      for (int pair = team; pair < n_pairs; pair += n_teams) {
        const int nab = min(K, p.n_clips - K * pair);
        for (int L = L0; L <= p.n_layers; ++L) {            // L = 0 is conv0
          for (int ab = 0; ab < nab; ++ab, ++ev) {
            mbar_wait(a_done + 8 * (ev & 3), (ev >> 2) & 1);
            fence_proxy_async_global();
            red_release_gpu_add(p.flags + K * pair + ab, 1);
          }
        }
      }
    }
  }

  fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// MFCC [n_clips][H][W] float32 -> conv0's A operand [n_clips][W][H+2][8] f16 = (hi, lo, hi, 0, 0, 0, 0, 0) with a
// zero row above and below (the time padding of conv0).  hi + lo carries 22 significant bits of the MFCC.
__global__ void __launch_bounds__(256) tc2_pack_x0_kernel(const float* __restrict__ mfcc, unsigned char* __restrict__ x0, int H, int W) {
  extern __shared__ float tile[];                      // [H][W + 1]
  const float* src = mfcc + static_cast<long long>(blockIdx.x) * H * W;
  for (int i = threadIdx.x; i < H * W; i += blockDim.x) tile[(i / W) * (W + 1) + i % W] = __ldg(src + i);
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(x0) + static_cast<long long>(blockIdx.x) * W * (H + 2);
  for (int i = threadIdx.x; i < W * (H + 2); i += blockDim.x) {
    const int w = i / (H + 2), r = i - w * (H + 2);
    uint4 o = make_uint4(0, 0, 0, 0);
    if (r >= 1 && r <= H) {
      const float x = fminf(fmaxf(tile[(r - 1) * (W + 1) + w], -60000.f), 60000.f);   // stays finite in f16
      const __half hi = __float2half_rn(x);
      const __half lo = __float2half_rn(x - __half2float(hi));
      const uint32_t uh = __half_as_ushort(hi), ul = __half_as_ushort(lo);
      o.x = uh | (ul << 16);
      o.y = uh;
    }
    dst[i] = o;
  }
}

// Pooled models: conv0 (1 -> C, 3x3, pad 1, bias-free) + ReLU + AvgPool(ph, pw) in FP32 (model.py:107-111), written as
// the persistent kernel's input image y0[group][w][cg][plane_rows][8] f16: S clips stacked along the time axis, one zero
// row between them, T2_PAD zero rows above every plane, constant-one in the two spare channels at real positions.
// One CTA per group; every 16-byte unit of the group's buffer is written (zeros included), so y0 needs no memset.
// Thread t serves channel group t / 64 with its 72 weights in registers; a pooled cell's (PH+2) x (PW+2) input patch is
// read from shared memory once and reused by its PH*PW conv positions.  PH = 0: pool window taken from ph, pw at run time.
template <int NCG, int PH, int PW>
__global__ void __launch_bounds__(NCG * 64) tc2_conv0_pool_kernel(const float* __restrict__ mfcc, const float* __restrict__ w0,
                                                                  unsigned char* __restrict__ y0, int n_clips, int H0, int W0, int ph,
                                                                  int pw, int Hc, int Wc, int S, int plane_rows, long long slot_stride) {
  constexpr int N = NCG * 8;
  extern __shared__ float xs[];                          // [S][(H0+2)][(W0+2)] zero-bordered MFCC images
  const int group = blockIdx.x, tid = threadIdx.x;
  const int Wp = W0 + 2, img = (H0 + 2) * Wp;
  const int n_here = static_cast<int>(min(static_cast<long long>(S), n_clips - static_cast<long long>(group) * S));
  for (int i = tid; i < S * img; i += NCG * 64) {
    const int sc = i / img, r = i - sc * img;
    const int y = r / Wp - 1, x = r % Wp - 1;
    const long long clip = static_cast<long long>(group) * S + sc;
    xs[i] = (sc < n_here && y >= 0 && y < H0 && x >= 0 && x < W0) ? __ldg(mfcc + (clip * H0 + y) * W0 + x) : 0.f;
  }
  const int cg = tid >> 6, t64 = tid & 63;
  float wr[9][8];
#pragma unroll
  for (int tap = 0; tap < 9; ++tap)
#pragma unroll
    for (int k = 0; k < 8; ++k) wr[tap][k] = __ldg(w0 + tap * N + cg * 8 + k);
  const int Hp1 = Hc + 1, Hk = S * Hp1 - 1;
  uint4* dst = reinterpret_cast<uint4*>(y0 + static_cast<long long>(group) * slot_stride);
  // every unit that is not a real position of a present clip is zero: padding rows, separators, missing clips, tail
  const int n_units = static_cast<int>(slot_stride / 16), n_plane_units = Wc * NCG * plane_rows;
  for (int i = tid; i < n_units; i += NCG * 64) {
    bool real = false;
    if (i < n_plane_units) {
      const int hk = i % plane_rows - T2_PAD;
      if (hk >= 0 && hk < Hk) {
        const int sc = hk / Hp1;
        real = (hk - sc * Hp1) < Hc && sc < n_here;
      }
    }
    if (!real) dst[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  const float inv = 1.f / static_cast<float>(ph * pw);
  const int rows = n_here * Hc;                          // real rows of one column of the group
  for (int j = t64; j < Wc * rows; j += 64) {
    const int w = j / rows, r = j - w * rows;
    const int sc = r / Hc, h = r - sc * Hc;
    const float* xi = xs + sc * img + (h * ph) * Wp + w * pw;     // top-left of the cell's input patch (bordered image)
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    if (PH > 0) {
      float patch[(PH > 0 ? PH : 1) + 2][(PW > 0 ? PW : 1) + 2];
#pragma unroll
      for (int y = 0; y < PH + 2; ++y)
#pragma unroll
        for (int x = 0; x < PW + 2; ++x) patch[y][x] = xi[y * Wp + x];
#pragma unroll
      for (int py = 0; py < PH; ++py)
#pragma unroll
        for (int px = 0; px < PW; ++px) {
          float t[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) t[k] = 0.f;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const float v = patch[py + tap / 3][px + tap % 3];
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] = fmaf(v, wr[tap][k], t[k]);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += fmaxf(t[k], 0.f);
        }
    } else {
      for (int py = 0; py < ph; ++py)
        for (int px = 0; px < pw; ++px) {
          float t[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) t[k] = 0.f;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const float v = xi[(py + tap / 3) * Wp + px + tap % 3];
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] = fmaf(v, wr[tap][k], t[k]);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += fmaxf(t[k], 0.f);
        }
    }
    uint4 o;
    o.x = pack_h2(acc[0] * inv, acc[1] * inv);
    o.y = pack_h2(acc[2] * inv, acc[3] * inv);
    o.z = pack_h2(acc[4] * inv, acc[5] * inv);
    o.w = pack_h2(acc[6] * inv, acc[7] * inv);
    if (cg == NCG - 1) o.w = 0x3C003C00u;                // the two constant-one channels
    dst[(static_cast<long long>(w) * NCG + cg) * plane_rows + T2_PAD + sc * Hp1 + h] = o;
  }
}

// D(16x8, f32) += A(16x16, f16, row) * B(16x8, f16, col): the warp-level tensor-core MMA (legacy mma.sync path; conv0 is
// 0.2 % of a model's work and its M = positions x N = 48 x K = 28 shape is far too small for a tcgen05 pipeline of its own)
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Pooled models, default pre-pass: conv0 + ReLU + AvgPool as an implicit GEMM on the tensor cores.  M = 16 pooled cells
// (one conv position of each per MMA), N = 8 channels per tile, K = 32 = 16 pairs:
//   pair m = 0..8  : A = (x_hi, x_lo) of tap m        B = (w_hi, w_hi) of tap m
//   pair m = 9..13 : A = (x_hi of taps 2(m-9), 2(m-9)+1)   B = (w_lo of the same two taps)      (tap 9 does not exist: 0)
// i.e. x*w_hi + x_hi*w_lo with x = x_hi + x_lo, w = w_hi + w_lo split into f16 halves and FP32 accumulation (~2^-21
// relative, like conv0 of the un-pooled models).  The shared-memory image holds one (hi, lo) pair per pixel, so a pair of
// the first kind is one 32-bit load.  ReLU and the pool sum are applied to the accumulator fragment per conv position.
template <int NCG>
__global__ void __launch_bounds__(256) tc2_conv0_pool_mma_kernel(const float* __restrict__ mfcc, const uint32_t* __restrict__ wb,
                                                                 unsigned char* __restrict__ y0, int n_clips, int H0, int W0, int ph,
                                                                 int pw, int Hc, int Wc, int S, int plane_rows, long long slot_stride) {
  extern __shared__ uint32_t xh[];                       // [S][(H0+2)][(W0+2)] zero-bordered images, (hi | lo << 16) per pixel
  const int group = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int Wp = W0 + 2, img = (H0 + 2) * Wp;
  const int n_here = static_cast<int>(min(static_cast<long long>(S), n_clips - static_cast<long long>(group) * S));
  auto split = [](float f) -> uint32_t {
    f = fminf(fmaxf(f, -60000.f), 60000.f);              // stays finite in f16
    const __half hi = __float2half_rn(f);
    const __half lo = __float2half_rn(f - __half2float(hi));
    return static_cast<uint32_t>(__half_as_ushort(hi)) | (static_cast<uint32_t>(__half_as_ushort(lo)) << 16);
  };
  const float* src = mfcc + static_cast<long long>(group) * S * H0 * W0;     // the group's clips are contiguous
  if ((W0 & 3) == 0 && ((static_cast<long long>(S) * H0 * W0) & 3) == 0) {
    // interior: 16-byte loads, four in flight per thread (the border and absent clips are zeroed below)
    const float4* src4 = reinterpret_cast<const float4*>(src);
    const int n4 = n_here * H0 * W0 / 4, hw = H0 * W0;
    for (int i0 = tid; i0 < n4; i0 += 4 * 256) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * 256 < n4) v[u] = __ldg(src4 + i0 + u * 256);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = (i0 + u * 256) * 4;
        if (e < n4 * 4) {
          const int sc = e / hw, rem = e - sc * hw, y = rem / W0, x = rem - y * W0;
          uint32_t* d = xh + sc * img + (y + 1) * Wp + x + 1;
          d[0] = split(v[u].x);
          d[1] = split(v[u].y);
          d[2] = split(v[u].z);
          d[3] = split(v[u].w);
        }
      }
    }
    for (int i = tid; i < S * img; i += 256) {
      const int sc = i / img, r = i - sc * img;
      const int y = r / Wp, x = r - y * Wp;
      if (sc >= n_here || y == 0 || y == H0 + 1 || x == 0 || x == W0 + 1) xh[i] = 0u;
    }
  } else {
    for (int i = tid; i < S * img; i += 256) {
      const int sc = i / img, r = i - sc * img;
      const int y = r / Wp - 1, x = r % Wp - 1;
      xh[i] = (sc < n_here && y >= 0 && y < H0 && x >= 0 && x < W0) ? split(__ldg(src + (static_cast<long long>(sc) * H0 + y) * W0 + x)) : 0u;
    }
  }
  const int Hp1 = Hc + 1, Hk = S * Hp1 - 1;
  uint4* dst = reinterpret_cast<uint4*>(y0 + static_cast<long long>(group) * slot_stride);
  // every unit that is not a real position of a present clip is zero: padding rows, separators, missing clips, tail
  const int n_units16 = static_cast<int>(slot_stride / 16), n_plane_units = Wc * NCG * plane_rows;
  for (int i = tid; i < n_units16; i += 256) {
    bool real = false;
    if (i < n_plane_units) {
      const int hk = i % plane_rows - T2_PAD;
      if (hk >= 0 && hk < Hk) {
        const int sc = hk / Hp1;
        real = (hk - sc * Hp1) < Hc && sc < n_here;
      }
    }
    if (!real) dst[i] = make_uint4(0, 0, 0, 0);
  }
  // this thread's four K pairs (q = lane % 4): m = q, q + 4 (one pixel each) and q + 8, q + 12 (generic: two pixels + masks)
  const int q = lane & 3, r0 = lane >> 2;
  const int off0 = (q / 3) * Wp + q % 3, off1 = ((q + 4) / 3) * Wp + (q + 4) % 3;
  int off2a, off2b, off3a, off3b;
  uint32_t sel2, mask2b, z2, sel3, mask3b, z3;
  {
    auto tap_off = [&](int t) { return (t / 3) * Wp + t % 3; };
    const int m2 = q + 8, m3 = q + 12;
    if (m2 == 8) { off2a = off2b = tap_off(8); sel2 = 0x3210u; mask2b = 0xFFFFFFFFu; }
    else { off2a = tap_off(2 * (m2 - 9)); off2b = tap_off(2 * (m2 - 9) + 1); sel2 = 0x5410u; mask2b = 0xFFFFFFFFu; }
    z2 = 0xFFFFFFFFu;
    if (m3 == 12) { off3a = tap_off(6); off3b = tap_off(7); sel3 = 0x5410u; mask3b = 0xFFFFFFFFu; z3 = 0xFFFFFFFFu; }
    else if (m3 == 13) { off3a = off3b = tap_off(8); sel3 = 0x5410u; mask3b = 0u; z3 = 0xFFFFFFFFu; }
    else { off3a = off3b = 0; sel3 = 0x3210u; mask3b = 0u; z3 = 0u; }
  }
  uint32_t bw[NCG][4];
#pragma unroll
  for (int cg = 0; cg < NCG; ++cg) {
    const uint32_t* wr = wb + (cg * 8 + r0) * 16;        // B fragment: n = lane / 4
    bw[cg][0] = __ldg(wr + q);
    bw[cg][1] = __ldg(wr + q + 4);
    bw[cg][2] = __ldg(wr + q + 8);
    bw[cg][3] = __ldg(wr + q + 12);
  }
  __syncthreads();
  const float inv = 1.f / static_cast<float>(ph * pw);
  const int n_cells = Wc * Hc, units_per_clip = (n_cells + 15) >> 4;
  for (int u = warp; u < n_here * units_per_clip; u += 8) {
    const int sc = u / units_per_clip, c_base = (u - sc * units_per_clip) << 4;
    const int ca = c_base + r0, cb = ca + 8;
    const bool va = ca < n_cells, vb = cb < n_cells;
    const int wa = va ? ca / Hc : 0, ha = va ? ca - wa * Hc : 0;
    const int wbq = vb ? cb / Hc : 0, hb = vb ? cb - wbq * Hc : 0;
    const uint32_t* pa = xh + sc * img + (ha * ph) * Wp + wa * pw;       // top-left pixel of the cell's first conv position
    const uint32_t* pb = xh + sc * img + (hb * ph) * Wp + wbq * pw;
    float acc[NCG][4];
#pragma unroll
    for (int cg = 0; cg < NCG; ++cg)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[cg][k] = 0.f;
    for (int py = 0; py < ph; ++py)
      for (int px = 0; px < pw; ++px) {
        const uint32_t* xa = pa + py * Wp + px;
        const uint32_t* xb = pb + py * Wp + px;
        const uint32_t a00 = xa[off0], a01 = xb[off0], a10 = xa[off1], a11 = xb[off1];
        const uint32_t a20 = __byte_perm(xa[off2a], xa[off2b] & mask2b, sel2) & z2;
        const uint32_t a21 = __byte_perm(xb[off2a], xb[off2b] & mask2b, sel2) & z2;
        const uint32_t a30 = __byte_perm(xa[off3a], xa[off3b] & mask3b, sel3) & z3;
        const uint32_t a31 = __byte_perm(xb[off3a], xb[off3b] & mask3b, sel3) & z3;
        float c[NCG][4];                                  // NCG independent accumulator chains per K step
#pragma unroll
        for (int cg = 0; cg < NCG; ++cg) {
          c[cg][0] = c[cg][1] = c[cg][2] = c[cg][3] = 0.f;
          mma16816(c[cg], a00, a01, a10, a11, bw[cg][0], bw[cg][1]);
        }
#pragma unroll
        for (int cg = 0; cg < NCG; ++cg) mma16816(c[cg], a20, a21, a30, a31, bw[cg][2], bw[cg][3]);
#pragma unroll
        for (int cg = 0; cg < NCG; ++cg)
#pragma unroll
          for (int k = 0; k < 4; ++k) acc[cg][k] += fmaxf(c[cg][k], 0.f);
      }
    // rows r0 / r0 + 8 of the tile = cells ca / cb; this lane holds channels 2q, 2q + 1 of every channel group
    unsigned char* oa = reinterpret_cast<unsigned char*>(dst) + ((static_cast<long long>(wa) * NCG) * plane_rows + T2_PAD + sc * Hp1 + ha) * 16 + q * 4;
    unsigned char* ob = reinterpret_cast<unsigned char*>(dst) + ((static_cast<long long>(wbq) * NCG) * plane_rows + T2_PAD + sc * Hp1 + hb) * 16 + q * 4;
#pragma unroll
    for (int cg = 0; cg < NCG; ++cg) {
      uint32_t o0 = pack_h2(acc[cg][0] * inv, acc[cg][1] * inv), o1 = pack_h2(acc[cg][2] * inv, acc[cg][3] * inv);
      if (cg == NCG - 1 && q == 3) o0 = o1 = 0x3C003C00u;      // the two constant-one channels
      if (va) *reinterpret_cast<uint32_t*>(oa + static_cast<long long>(cg) * plane_rows * 16) = o0;
      if (vb) *reinterpret_cast<uint32_t*>(ob + static_cast<long long>(cg) * plane_rows * 16) = o1;
    }
  }
}

// Pooled models, default pre-pass since r04: conv0 + ReLU + AvgPool on tcgen05.  One CTA walks groups of S clips; per clip
//   * 4 builder warps turn the MFCCs into conv0's operand image in shared memory (double-buffered): per coefficient
//     plane w and row r a 16-byte unit (x_hi, x_lo, x_hi, 0, x'_hi, x'_lo, x'_hi, 0) of frames r and r' = r + 1, with zero
//     frames above and below the clip;
//   * one warp issues ONE M128 x N(3 NB) x K16 MMA per INPUT column c: the A descriptor's second K chunk sits two rows
//     (32 bytes) after the first, so the chunks carry the time taps kh = (0, 1) and (2, -), and the three coefficient taps kw
//     are stacked along N against B = (w_hi, w_hi, w_lo) per tap: block kw lands on OUTPUT column c - (kw - 1).  Output
//     columns sit in adjacent NB-column blocks of a ring of three TMEM buffers (one per pooled column = pw outputs), so the
//     three contributions to an output add up inside the accumulator ("ring accumulate", as in tc2_res_kernel); an MMA
//     that would straddle the ring's end is issued in two pieces;
//   * 8 epilogue warps (two per TMEM lane quarter, half the channels each) read a finished buffer, hand it back ZEROED
//     (every MMA accumulates), apply the ReLU per conv position, add the pw columns, add the ph time rows with a
//     halving exchange of lane shuffles and store f16 into the persistent kernel's stacked input image y0 (same bytes as
//     tc2_conv0_pool_mma_kernel writes, zero units included).
// x*w is (x_hi + x_lo)*w_hi + x_hi*w_lo with FP32 accumulation, like both other conv0 formulations.
constexpr int T2C0_NI = 1;            // MMA-issuing warps (more than one makes the accumulation order, hence the last bit, vary run to run)
constexpr int T2C0_E0 = T2C0_NI + 4;  // first epilogue warp (T2C0_E0 % 4 may be anything: the TMEM lane quarter is warp % 4)
#ifndef T2C0_NSET
#define T2C0_NSET 1
#endif
constexpr int T2C0_THREADS = (T2C0_E0 + 8 * T2C0_NSET) * 32;   // issuer, 4 builders, T2C0_NSET sets of 8 epilogue warps (alternate pooled columns)
constexpr int T2C0_PADROWS = 32;      // rows past the last plane that the 128-row A tiles (+3 rows of taps) may read
constexpr int T2C0_NBUF = 3;
constexpr int T2C0_W0 = 40;           // MFCC coefficients per frame (compile-time)
constexpr int T2C0_TAIL = 2 * 3 * 48 * 16 + 256;   // B tile (largest NB) + barriers + TMEM slot, after the two images


// One of the T2C0_NI issuing warps of tc2_conv0_pool_tc_kernel: input columns c with c % T2C0_NI == WI.  Pooled column j of
// every clip uses ring buffer j % 3, so the per-column schedule (which outputs an input column touches, where they sit in
// TMEM, which buffer this warp touches first / last) unrolls into constants.  A warp waits for a buffer before ITS first MMA
// into it and commits to the buffer's "full" barrier after ITS last one (the barrier counts T2C0_NI arrivals).
template <int NCG, int PW, int WI>
__device__ __forceinline__ void c0_issue(int n_groups, int n_clips, int S, int Hi, uint32_t img_a, int img_bytes, uint32_t wt_a,
                                         uint32_t tmem_base, uint32_t b_img_full, uint32_t b_img_empty, uint32_t b_grp_full,
                                         uint32_t b_grp_empty, int mode) {
  constexpr int NB = NCG * 8, W0 = T2C0_W0, Wc = W0 / PW, n_out = Wc * PW, BUF = PW * NB, NI = T2C0_NI;
  uint32_t n = 0, uses[T2C0_NBUF] = {0, 0, 0};                   // clips started; uses of each ring buffer so far
  uint64_t bdesc[3];
#pragma unroll
  for (int blk = 0; blk < 3; ++blk) bdesc[blk] = make_desc(wt_a + blk * NB * 16, 3 * NB * 16, 128);
  for (int g = blockIdx.x; g < n_groups; g += gridDim.x) {
    const int n_here = min(S, n_clips - g * S);
    for (int sc = 0; sc < n_here; ++sc, ++n) {
      const uint32_t ib = n & 1;
      mbar_wait(b_img_full + 8 * ib, (n >> 1) & 1);
      fence_after();
      const uint64_t a_base = make_desc(img_a + ib * img_bytes, 32, 128);
      const uint32_t plane16 = static_cast<uint32_t>(Hi);         // plane stride in 16-byte units
#pragma unroll
      for (int c = WI; c < W0; c += NI) {
        const int o_lo = c - 1 < 0 ? 0 : c - 1, o_hi = c + 1 > n_out - 1 ? n_out - 1 : c + 1;
        if (o_lo > o_hi) continue;
        const int g_lo = o_lo / PW, g_hi = o_hi / PW;
        // this warp's first MMA into a pooled column's buffer: it must have been drained (and zeroed) by the epilogue.  The
        // inputs of pooled column gg are columns max(0, gg*PW - 1) .. min(W0 - 1, gg*PW + PW).
#pragma unroll
        for (int gg = g_lo; gg <= g_hi; ++gg) {
          const int first_in = gg * PW - 1 < 0 ? 0 : gg * PW - 1;
          const int mine = first_in + ((WI - first_in) % NI + NI) % NI;                  // first column >= first_in of this warp
          if (c == mine) {
            mbar_wait(b_grp_empty + 8 * (gg % T2C0_NBUF), (uses[gg % T2C0_NBUF] & 1) ^ 1);
            ++uses[gg % T2C0_NBUF];
            fence_after();
          }
        }
        if (elect_one()) {
          const uint64_t adesc = a_base + static_cast<uint64_t>(c * plane16);
          const int b_lo = g_lo % T2C0_NBUF, b_hi = g_hi % T2C0_NBUF;
          const int split = (g_lo != g_hi && b_hi != b_lo + 1) ? g_hi * PW : o_hi + 1;    // first output of the second piece
#pragma unroll
          for (int piece = 0; piece < 2; ++piece) {
            const int oa = piece == 0 ? o_lo : split, ob = piece == 0 ? (o_hi < split - 1 ? o_hi : split - 1) : o_hi;
            if (oa > ob || (mode & 4)) continue;
            const int ga = oa / PW;
            mma_f16_ss(tmem_base + (ga % T2C0_NBUF) * BUF + (oa - ga * PW) * NB, adesc, bdesc[oa - (c - 1)],
                       make_idesc(128, (ob - oa + 1) * NB, 0), 1);
          }
          // this warp's last MMA into a pooled column's buffer
#pragma unroll
          for (int gg = g_lo; gg <= g_hi; ++gg) {
            const int last_in = gg == Wc - 1 ? (W0 - 1 < n_out ? W0 - 1 : n_out) : gg * PW + PW;
            const int mine = last_in - ((last_in - WI) % NI + NI) % NI;                  // last column <= last_in of this warp
            if (c == mine) commit(b_grp_full + 8 * (gg % T2C0_NBUF));
          }
          if (c + NI >= W0) commit(b_img_empty + 8 * ib);          // this warp's last column: its reads of the image are done
        }
        __syncwarp();
      }
    }
  }
}

template <int NCG, int PW>
__global__ void __launch_bounds__(T2C0_THREADS, 1) tc2_conv0_pool_tc_kernel(const float* __restrict__ mfcc, const unsigned char* __restrict__ wt,
                                                                            unsigned char* __restrict__ y0, int n_clips, int H0, int ph,
                                                                            int Hc, int S, int plane_rows, long long slot_stride,
                                                                            int img_bytes, int mode) {
  constexpr int W0 = T2C0_W0, pw = PW, Wc = W0 / PW;
  // mode (HONK_TC2_C0_MODE, timing experiments only - results are garbage): bit 0 builders skip their loads and stores,
  // bit 1 the epilogue skips its pooling and stores, bit 2 the issuer skips the MMAs, bit 3 no zeroing of y0's pad units
  constexpr int NB = NCG * 8, HB = NB / 2, TILE = 2 * 3 * NB * 16;
  extern __shared__ __align__(128) unsigned char c0_smem[];
  unsigned char* wts = c0_smem + 2 * img_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(wts + TILE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4 + 2 * T2C0_NBUF);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Hi = H0 + 2, plane_bytes = Hi * 16, HW = H0 * W0;
  const uint32_t img_a = smem_u32(c0_smem), wt_a = smem_u32(wts);
  const uint32_t b_img_full = smem_u32(bars + 0), b_img_empty = smem_u32(bars + 2), b_grp_full = smem_u32(bars + 4),
                 b_grp_empty = smem_u32(bars + 4 + T2C0_NBUF);
  for (int i = tid; i < 2 * img_bytes / 16; i += T2C0_THREADS) reinterpret_cast<uint4*>(c0_smem)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < TILE / 16; i += T2C0_THREADS) reinterpret_cast<uint4*>(wts)[i] = __ldg(reinterpret_cast<const uint4*>(wt) + i);
  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(b_img_full + 8 * b, 128);
      mbar_init(b_img_empty + 8 * b, T2C0_NI);
    }
    for (int b = 0; b < T2C0_NBUF; ++b) {
      mbar_init(b_grp_full + 8 * b, T2C0_NI);
      mbar_init(b_grp_empty + 8 * b, 8);
    }
    fence_mbar_init();
  }
  fence_async_smem();
  if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 512);
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tmem_base = *tmem_slot;
  static_assert(Wc * PW >= W0 - 1 && PW >= 2 && T2C0_NBUF * PW * NB <= 512, "schedule assumes every input column feeds an output");
  const int n_groups = (n_clips + S - 1) / S, Hp1 = Hc + 1;
  constexpr int buf_cols = PW * NB;
  if (warp >= T2C0_E0) {                                        // all accumulators start at zero: every MMA accumulates
    const int lq = warp & 3, ch = ((warp - T2C0_E0) & 7) >> 2;
    for (int i = 0; i < T2C0_NBUF * PW; ++i)
      tmem_zero_cols<HB>(tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + i * NB + ch * HB);
    tmem_st_wait();
  }
  fence_before();
  __syncthreads();
  fence_after();

  if (warp < T2C0_NI) {
    // ---- MMA issuers: T2C0_NI warps take alternate input columns (all MMAs accumulate, so their order does not matter) ------
    if (T2C0_NI == 1 || warp == 0)
      c0_issue<NCG, PW, 0>(n_groups, n_clips, S, Hi, img_a, img_bytes, wt_a, tmem_base, b_img_full, b_img_empty, b_grp_full, b_grp_empty, mode);
    else
      c0_issue<NCG, PW, T2C0_NI - 1>(n_groups, n_clips, S, Hi, img_a, img_bytes, wt_a, tmem_base, b_img_full, b_img_empty, b_grp_full, b_grp_empty, mode);
  } else if (warp < T2C0_NI + 4) {
    // ---- operand-image builders (and the zero units of the group's output buffer) -----------------------------------
    const int bt = tid - 32 * T2C0_NI;
    uint32_t n = 0;
    for (int g = blockIdx.x; g < n_groups; g += gridDim.x) {
      const int n_here = min(S, n_clips - g * S);
      if (!(mode & 8)) {
        // zero units of the group's buffer: rows that are not a real position of a present clip (same test for every plane,
        // made once per lane for its rows lane, lane + 32, ...), and the slot's tail
        const int Hk = S * Hp1 - 1, bw = warp - T2C0_NI;
        uint4* dst = reinterpret_cast<uint4*>(y0 + static_cast<long long>(g) * slot_stride);
        const int n_units16 = static_cast<int>(slot_stride / 16), n_planes = Wc * NCG;
        uint32_t zero_rows = 0;
        for (int k = 0; k * 32 + lane < plane_rows; ++k) {
          const int hk = k * 32 + lane - T2_PAD;
          bool real = false;
          if (hk >= 0 && hk < Hk) {
            const int sc = hk / Hp1;
            real = (hk - sc * Hp1) < Hc && sc < n_here;
          }
          if (!real) zero_rows |= 1u << k;
        }
        for (int pl = bw; pl < n_planes; pl += 4)
          for (int k = 0; k * 32 + lane < plane_rows; ++k)
            if ((zero_rows >> k) & 1) dst[pl * plane_rows + k * 32 + lane] = make_uint4(0, 0, 0, 0);
        for (int i = n_planes * plane_rows + bt; i < n_units16; i += 128) dst[i] = make_uint4(0, 0, 0, 0);
      }
      for (int sc = 0; sc < n_here; ++sc, ++n) {
        const uint32_t ib = n & 1;
        const float* src = mfcc + (static_cast<long long>(g) * S + sc) * HW;
        unsigned char* img = c0_smem + ib * img_bytes;
        bool waited = false;
        if (mode & 1) {
          mbar_wait(b_img_empty + 8 * ib, ((n >> 1) & 1) ^ 1);
          waited = true;
        }
        for (int e0 = bt; e0 < HW && !(mode & 1); e0 += 128 * 16) {
          float v[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) v[u] = (e0 + u * 128 < HW) ? __ldg(src + e0 + u * 128) : 0.f;
          if (!waited) {                                         // the loads are in flight while the buffer drains
            mbar_wait(b_img_empty + 8 * ib, ((n >> 1) & 1) ^ 1);
            waited = true;
          }
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const int e = e0 + u * 128;
            if (e < HW) {
              const int h = e / W0, w = e - h * W0;
              const float x = fminf(fmaxf(v[u], -60000.f), 60000.f);                 // stays finite in f16
              const __half hi = __float2half_rn(x);
              const __half lo = __float2half_rn(x - __half2float(hi));
              const uint32_t uh = __half_as_ushort(hi), ul = __half_as_ushort(lo);
              const uint2 val = make_uint2(uh | (ul << 16), uh);
              unsigned char* unit = img + (w * Hi + h + 1) * 16;                     // frame h is image row h + 1
              *reinterpret_cast<uint2*>(unit) = val;                                 // as frame r of its own row
              *reinterpret_cast<uint2*>(unit - 16 + 8) = val;                        // as frame r + 1 of the row above
            }
          }
        }
        fence_async_smem();
        mbar_arrive(b_img_full + 8 * ib);
      }
    }
  } else {
    // ---- epilogue: ReLU per conv position, sum of the pool window, f16 store ---------------------------------------------
    // Two sets of 8 warps take alternate pooled columns.  A warp owns one TMEM lane quarter (32 frames) and half the channels:
    // it adds relu() of the PW conv columns per frame, hands the buffer back zeroed, then pools ph frames through a private
    // shared-memory tile (row stride HB + 4 floats: conflict-free 16-byte stores) - lane l sums the ph rows of cell l / nq for
    // its share l % nq of the channels and stores them.
    const int ew = (warp - T2C0_E0) & 7, set = (warp - T2C0_E0) >> 3, lq = warp & 3, ch = ew >> 2;   // TMEM lane quarter = warp id % 4
    constexpr int TS = HB + 4;                                   // tile row stride in floats
    float* tile = reinterpret_cast<float*>(c0_smem + 2 * img_bytes + T2C0_TAIL) + (warp - T2C0_E0) * 32 * TS;
    // pooling role of this lane (ph = 4: 8 cells x 4 channel quarters; ph = 2: 16 cells x 2 channel halves)
    const int nq = ph == 4 ? 4 : 2, cell = lane / nq, part = lane - cell * nq, PC = HB / nq;      // PC channels per lane
    const int hp = (lq * 32) / ph + cell;
    const bool row_ok = hp < Hc && (ph == 4 || ph == 2);
    const float inv = 1.f / static_cast<float>(ph * PW);
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(lq * 32) << 16) + ch * HB;
    const int chn0 = ch * HB + part * PC;                        // first channel this lane stores
    uint32_t uses[T2C0_NBUF] = {0, 0, 0}, turn = 0;
    for (int g = blockIdx.x; g < n_groups; g += gridDim.x) {
      const int n_here = min(S, n_clips - g * S);
      for (int sc = 0; sc < n_here; ++sc) {
        // this lane's first 4-byte store of pooled column 0: unit (j * NCG + chn / 8) * plane_rows + T2_PAD + sc * Hp1 + hp
        unsigned char* out = y0 + static_cast<long long>(g) * slot_stride + (static_cast<long long>(T2_PAD + sc * Hp1 + hp)) * 16;
#pragma unroll 3
        for (int j = 0; j < Wc; ++j, ++turn) {
          const uint32_t b = j % T2C0_NBUF;
          const uint32_t phase = (b == 0 ? uses[0] : (b == 1 ? uses[1] : uses[2])) & 1;
          if (b == 0) ++uses[0]; else if (b == 1) ++uses[1]; else ++uses[2];
          if (T2C0_NSET > 1 && (turn % T2C0_NSET) != static_cast<uint32_t>(set)) continue;      // another set's pooled column
          mbar_wait(b_grp_full + 8 * b, phase);
          fence_after();
          const uint32_t taddr = t_lane + b * buf_cols;
          float acc[HB];
          uint32_t v[PW][HB];
#pragma unroll
          for (int c = 0; c < PW; ++c) tmem_ld_cols<HB>(taddr + c * NB, v[c]);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < PW; ++c) tmem_zero_cols<HB>(taddr + c * NB);
#pragma unroll
          for (int k = 0; k < HB; ++k) {
            float a = fmaxf(__uint_as_float(v[0][k]), 0.f);
#pragma unroll
            for (int c = 1; c < PW; ++c) a += fmaxf(__uint_as_float(v[c][k]), 0.f);
            acc[k] = a;
          }
          tmem_st_wait();
          fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(b_grp_empty + 8 * b);           // the buffer goes back zeroed: every MMA accumulates
          if (mode & 2) continue;
#pragma unroll
          for (int k = 0; k < HB; k += 4) *reinterpret_cast<float4*>(tile + lane * TS + k) = make_float4(acc[k], acc[k + 1], acc[k + 2], acc[k + 3]);
          __syncwarp();
          if (row_ok) {
            const float* src = tile + cell * ph * TS + part * PC;
            if (ph == 4) {
#pragma unroll
              for (int k = 0; k < HB / 4; k += 2) {
                const float2 r0 = *reinterpret_cast<const float2*>(src + k), r1 = *reinterpret_cast<const float2*>(src + TS + k),
                             r2 = *reinterpret_cast<const float2*>(src + 2 * TS + k), r3 = *reinterpret_cast<const float2*>(src + 3 * TS + k);
                const int chn = chn0 + k;
                uint32_t o = pack_h2(((r0.x + r1.x) + (r2.x + r3.x)) * inv, ((r0.y + r1.y) + (r2.y + r3.y)) * inv);
                if (chn >= NB - 2) o = 0x3C003C00u;                                  // the two constant-one channels
                *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(j * NCG + (chn >> 3)) * plane_rows) * 16 + (chn & 7) * 2) = o;
              }
            } else {
#pragma unroll
              for (int k = 0; k < HB / 2; k += 2) {
                const float2 r0 = *reinterpret_cast<const float2*>(src + k), r1 = *reinterpret_cast<const float2*>(src + TS + k);
                const int chn = chn0 + k;
                uint32_t o = pack_h2((r0.x + r1.x) * inv, (r0.y + r1.y) * inv);
                if (chn >= NB - 2) o = 0x3C003C00u;
                *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(j * NCG + (chn >> 3)) * plane_rows) * 16 + (chn & 7) * 2) = o;
              }
            }
          }
          __syncwarp();                                              // the tile is free for this warp's next pooled column
        }
      }
    }
  }
  fence_before();
  __syncthreads();
  if (warp == 0) {
    fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// Zeroes exactly the 16-byte units of the activation arena that no layer ever stores and that must read as zero: the
// T2_PAD rows above every channel plane (time padding), the S - 1 separator rows of a stacked group, and each slot's tail.
__global__ void __launch_bounds__(256) tc2_zero_pads_kernel(unsigned char* __restrict__ act, long long slot_stride, int n_slots2,
                                                            int planes_per_slot, int plane_rows, int S, int Hc, int tail_units) {
  const int per_plane = T2_PAD + S - 1;
  const long long per_slot = static_cast<long long>(planes_per_slot) * per_plane + tail_units;
  const long long total = per_slot * n_slots2;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long slot = i / per_slot, r = i - slot * per_slot;
    long long unit;
    if (r < static_cast<long long>(planes_per_slot) * per_plane) {
      const long long pl = r / per_plane;
      const int k = static_cast<int>(r - pl * per_plane);
      const int row = k < T2_PAD ? k : T2_PAD + (k - T2_PAD) * (Hc + 1) + Hc;
      unit = pl * plane_rows + row;
    } else {
      unit = static_cast<long long>(planes_per_slot) * plane_rows + (r - static_cast<long long>(planes_per_slot) * per_plane);
    }
    *reinterpret_cast<uint4*>(act + slot * slot_stride + unit * 16) = make_uint4(0, 0, 0, 0);
  }
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
  long long slot_stride = 0, x0_clip_stride = 0;
  int max_teams = 37;
  int max_resident = 148;         // co-resident CTAs of tc2_res_kernel on this device (occupancy x SMs)
  int G = T2_G;                   // CTAs per team
  int S = 1, Hc = 101;            // pooled models: S clips of Hc rows stacked along time per kernel "clip" (H = S*(Hc+1)-1)
  int pool_h = 0, pool_w = 0, H0 = 101, W0 = 40;
  float* d_w0 = nullptr;          // pooled models: conv0 weights [9][NB] float (FP32 variant of the conv0 + pool pre-pass)
  uint32_t* d_w0b = nullptr;      // pooled models: conv0 weights as mma.sync B pairs [NB][16] (tc2_conv0_pool_mma_kernel)
  bool conv0_f32 = false;         // HONK_TC2_CONV0_F32=1: FP32 CUDA-core pre-pass instead of the tensor-core one
  __half* d_w0t = nullptr;        // pooled models: conv0 weights as six tcgen05 B tiles [kw][kh pair | kh 2][2 K chunks][NB][8 halves]
  bool conv0_tc = false;          // tcgen05 pre-pass (tc2_conv0_pool_tc_kernel); HONK_TC2_CONV0_MMA=1 keeps the mma.sync one
  int c0_img_bytes = 0, c0_smem = 0;
  int K = 2;                      // clips per team in flight: 37 teams x 2 x 2 buffers x 449 KB = 66 MB of L2 (K = 3: 100 MB starts
                                  // evicting dirty activations to HBM, and the extra DRAM power costs SM clock under the power cap)
  std::vector<T2Layer> layers;    // [0] = conv0, [1..n_layers] = the C->C layers
  std::vector<T2Layer> layers_lat; // same layers cut for teams of T2_G CTAs: request-sized batches (latency mode)
  int n_sms = 148;
  bool tuned_by_env = false;      // HONK_TC2_G / HONK_TC2_K given: no automatic latency mode
  unsigned char* d_wpack = nullptr;
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
bool plan_segments(T2Layer& ly, int W, int d, int G) {
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
  if (n_chain < G) {
    const int pieces = G / n_chain;                // n_chain is 1 or 2 here
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
    for (int k = 1; k < G; ++k)
      if (load[k] < load[best]) best = k;
    if (!add(best, c.second, 0, c.first, c.first)) return false;
    load[best] += c.first;
  }
  return true;
}

template <typename Kern>
int set_smem(Kern k, int bytes) {
  HONK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return HONK_OK;
}

}  // namespace

int res_tc2_plan_create(honk_model* m, ResTc2Plan** out) {
  *out = nullptr;
  const honk_model_config& c = m->cfg;
  const int C = c.n_feature_maps;
  if (c.family != HONK_FAMILY_RES || C > 46) return HONK_OK;   // two spare channels carry the mean shift
  const bool pooled = c.res_pool[0] > 0;
  // Pooled models (res8: 25 x 13, res26: 50 x 20 maps): S clips are stacked along the time axis with one zero row
  // between them, so one 128-row MMA tile carries S clips (res26: 50+1+50 = 101 rows, res8: 4*26-1 = 103 rows) and the
  // kernel sees an un-dilated "clip" of H = S*(Hc+1)-1 rows whose separator rows are never stored.
  const int S = pooled ? std::max(1, std::min(8, 113 / (m->Hr + 1))) : 1;
  const int Hk = pooled ? S * (m->Hr + 1) - 1 : m->H0, Wk = pooled ? m->Wr : m->W0;
  int G = T2_G;
  if (pooled) {
    // measured (scripts/pooled_sweep.py, profiles/r03_pooled_sweep.json): res8's 13 columns are too few to share (one CTA
    // per group, no cross-CTA halo), res26's 20 columns run best on pairs of CTAs
    G = Wk <= 16 ? 1 : 2;
    if (const char* e = getenv("HONK_TC2_G")) {
      const int g = atoi(e);
      if (g == 1 || g == 2 || g == 4) G = g;
    }
    if (c.use_dilation && c.n_layers > 3) return HONK_OK;   // no reference config pools and dilates; one separator row = d 1
  }
  if (Hk > 112 || Hk < 8 || Wk < G || Wk > 1024 || c.n_layers < 1 || c.n_layers > T2_MAXL) return HONK_OK;
  for (int i = 1; i <= c.n_layers; ++i)
    if ((c.use_dilation ? (1 << ((i - 1) / 3)) : 1) > T2_PAD) return HONK_OK;
  std::unique_ptr<ResTc2Plan, void (*)(ResTc2Plan*)> p(new ResTc2Plan(), res_tc2_plan_free);
  p->C = C;
  p->ncg = C <= 30 ? 4 : 6;
  p->NB = p->ncg * 8;
  p->H = Hk;
  p->W = Wk;
  p->G = G;
  p->S = S;
  p->Hc = pooled ? m->Hr : m->H0;
  p->pool_h = pooled ? c.res_pool[0] : 0;
  p->pool_w = pooled ? c.res_pool[1] : 0;
  p->H0 = m->H0;
  p->W0 = m->W0;
  p->n_layers = c.n_layers;
  p->n_labels = c.n_labels;
  p->plane_rows = p->H + T2_PAD;
  p->col_bytes = p->ncg * p->plane_rows * 16;
  p->slot_stride = align_up(static_cast<int64_t>(p->W) * p->col_bytes + T2_PAD * 16, 256);
  p->x0_clip_stride = static_cast<long long>(p->W) * (p->H + 2) * 16;
  // a stage holds planes 0..ncg-1 of one column; the kh taps over-read up to 127 + 2*16 rows of the last plane
  p->stage_stride = static_cast<int>(align_up((p->ncg - 1) * p->plane_rows * 16 + (128 + 2 * T2_PAD) * 16, 128));
  const int tile_bytes = 6 * p->NB * 16;                       // one (kh, kc) B tile: 2 K halves x 3NB rows x 8 halves
  const int w0_bytes = pooled ? 0 : 3 * tile_bytes;
  p->wbytes = 3 * (p->ncg / 2) * tile_bytes;
  p->smem_bytes = T2_NSTAGE * p->stage_stride + 2 * p->wbytes;
  int max_smem = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
  if (p->smem_bytes + 2048 > max_smem) return HONK_OK;
  p->max_teams = std::max(1, m->n_sms / G);
  p->n_sms = m->n_sms;
  p->tuned_by_env = getenv("HONK_TC2_K") || getenv("HONK_TC2_G");
  if (const char* e = getenv("HONK_TC2_K")) p->K = std::max(1, std::min(8, atoi(e)));
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
  std::vector<T2Layer> layers(c.n_layers + 1);
  std::vector<unsigned char> wpack(static_cast<size_t>(w0_bytes) + static_cast<size_t>(c.n_layers) * p->wbytes, 0);
  if (pooled) {
    // conv0 + ReLU + AvgPool run before the persistent kernel, in FP32 (tc2_conv0_pool_kernel): weights [tap][NB]
    memset(&layers[0], 0, sizeof(T2Layer));
    const HostTensor& w = m->host.at("conv0.weight");
    std::vector<float> w0(9 * N, 0.f);
    for (int co = 0; co < C; ++co)
      for (int tap = 0; tap < 9; ++tap) w0[tap * N + co] = w.data[co * 9 + tap];
    HONK_TRY(up(p.get(), w0, &p->d_w0));
    // the same weights as K pairs of the mma.sync formulation (see tc2_conv0_pool_mma_kernel)
    std::vector<uint32_t> wb(static_cast<size_t>(N) * 16, 0u);
    for (int co = 0; co < C; ++co) {
      unsigned short hi[10] = {0}, lo[10] = {0};
      for (int tap = 0; tap < 9; ++tap) {
        const float v = w.data[co * 9 + tap];
        const __half h = __float2half(v);
        const __half l = __float2half(v - __half2float(h));
        hi[tap] = __half_as_ushort(h);
        lo[tap] = __half_as_ushort(l);
      }
      for (int mm = 0; mm < 9; ++mm) wb[co * 16 + mm] = static_cast<uint32_t>(hi[mm]) | (static_cast<uint32_t>(hi[mm]) << 16);
      for (int mm = 9; mm < 14; ++mm) wb[co * 16 + mm] = static_cast<uint32_t>(lo[2 * (mm - 9)]) | (static_cast<uint32_t>(lo[2 * (mm - 9) + 1]) << 16);
    }
    HONK_TRY(up(p.get(), wb, &p->d_w0b));
    {
      // tcgen05 formulation: ONE B tile [K chunk][row = blk * NB + co][8 halves], blk <-> kw = 2 - blk (block blk of input
      // column c lands on output column c - 1 + blk); chunk 0 = time taps kh 0 | 1, chunk 1 = kh 2 | none; a tap's four
      // halves are (w_hi, w_hi, w_lo, 0) against the operand image's (x_hi, x_lo, x_hi, 0)
      std::vector<__half> wt6(static_cast<size_t>(2) * 3 * N * 8, __float2half(0.f));
      for (int chunk = 0; chunk < 2; ++chunk)
        for (int half = 0; half < 2; ++half) {
          const int kh = chunk * 2 + half;
          if (kh > 2) continue;
          for (int blk = 0; blk < 3; ++blk) {
            const int kw = 2 - blk;
            for (int co = 0; co < C; ++co) {
              const float v = w.data[co * 9 + kh * 3 + kw];
              const __half hi = __float2half(v);
              const __half lo = __float2half(v - __half2float(hi));
              __half* row = wt6.data() + ((static_cast<size_t>(chunk) * 3 * N) + blk * N + co) * 8 + half * 4;
              row[0] = hi;
              row[1] = hi;
              row[2] = lo;
            }
          }
        }
      HONK_TRY(up(p.get(), wt6, &p->d_w0t));
      p->c0_img_bytes = static_cast<int>(align_up((static_cast<int64_t>(p->W0) * (p->H0 + 2) + T2C0_PADROWS) * 16, 128));
      p->c0_smem = 2 * p->c0_img_bytes + T2C0_TAIL + 8 * T2C0_NSET * 32 * (N / 2 + 4) * 4;
      const bool pow2 = (p->pool_h & (p->pool_h - 1)) == 0;
      p->conv0_tc = (p->pool_h == 2 || p->pool_h == 4) && (p->pool_w == 2 || p->pool_w == 3) && p->W0 == T2C0_W0 && p->W == p->W0 / p->pool_w && p->H0 <= 126 && p->c0_smem + 1024 <= max_smem &&
                    m->Hr * p->pool_h <= 128 && !(getenv("HONK_TC2_CONV0_MMA") && atoi(getenv("HONK_TC2_CONV0_MMA")) != 0);
    }
    if (const char* e = getenv("HONK_TC2_CONV0_F32")) p->conv0_f32 = atoi(e) != 0;
  } else {
    // conv0 (bias-free, model.py:107): B tile kh = [K half][n = blk*NB + co][8]; input channels 0..2 = (x_hi, x_lo, x_hi)
    // meet (w_hi, w_hi, w_lo); everything else (the whole second K half included) is zero.
    T2Layer& ly = layers[0];
    memset(&ly, 0, sizeof(ly));
    ly.d = 1;
    ly.flags = 4;
    ly.w_off = 0;
    ly.w_bytes = w0_bytes;
    if (!plan_segments(ly, p->W, 1, G)) return HONK_OK;
    const HostTensor& w = m->host.at("conv0.weight");
    __half* hw = reinterpret_cast<__half*>(wpack.data());
    for (int kh = 0; kh < 3; ++kh)
      for (int blk = 0; blk < 3; ++blk) {
        const int kw = 2 - blk;
        for (int co = 0; co < C; ++co) {
          const float v = w.data[co * 9 + kh * 3 + kw];
          const __half hi = __float2half(v);
          const __half lo = __float2half(v - __half2float(hi));
          __half* row = hw + ((static_cast<size_t>(kh) * 2 + 0) * (3 * N) + blk * N + co) * 8;
          row[0] = hi;
          row[1] = hi;
          row[2] = lo;
        }
      }
  }
  for (int L = 1; L <= c.n_layers; ++L) {
    T2Layer& ly = layers[L];
    memset(&ly, 0, sizeof(ly));
    ly.d = c.use_dilation ? (1 << ((L - 1) / 3)) : 1;
    const bool even = (L % 2 == 0), last = (L == c.n_layers);
    ly.flags = (even ? 1 : 0) | (last ? 2 : 0) | ((pooled && L == 1) ? 8 : 0) | ((pooled && L == 2) ? 16 : 0);
    ly.w_off = w0_bytes + (L - 1) * p->wbytes;
    ly.w_bytes = p->wbytes;
    if (!plan_segments(ly, p->W, ly.d, G)) return HONK_OK;
    // B operand: [i = kh*NKC + kc][K half][n = blk*NB + co][ci % 8], blk <-> kw = 2 - blk.
    //   ci <  C       : W[co][ci][kh][kw] * scale_{L-1}[ci]                 (BatchNorm scale of the producer)
    //   ci = NB-1/NB-2: hi / lo halves of  -sum_ci W*scale*mean_{L-1}[ci]   per TAP; these input channels hold
    //                   1.0 at real positions and 0 in the padding, so only in-range taps subtract the mean,
    //                   exactly like zero padding applied after bn_{L-1} (model.py:112-118)
    const HostTensor& w = m->host.at("conv" + std::to_string(L) + ".weight");
    __half* hw = reinterpret_cast<__half*>(wpack.data() + ly.w_off);
    auto slot = [&](int kh, int blk, int co, int ci) -> __half& {
      const int kc = ci / 16, khalf = (ci % 16) / 8, e = ci % 8;
      return hw[((static_cast<size_t>(kh * NKC + kc) * 2 + khalf) * (3 * N) + blk * N + co) * 8 + e];
    };
    for (int kh = 0; kh < 3; ++kh)
      for (int blk = 0; blk < 3; ++blk) {
        const int kw = 2 - blk;
        for (int co = 0; co < C; ++co) {
          double bias = 0.0;
          for (int ci = 0; ci < C; ++ci) {
            const float v = w.data[((static_cast<size_t>(co) * C + ci) * 3 + kh) * 3 + kw] * scale[L - 1][ci];
            const __half hv = __float2half(v);
            slot(kh, blk, co, ci) = hv;
            bias -= static_cast<double>(__half2float(hv)) * mean[L - 1][ci];   // mean of what the MMA really multiplies
          }
          const __half hi = __float2half(static_cast<float>(bias));
          const __half lo = __float2half(static_cast<float>(bias - static_cast<double>(__half2float(hi))));
          slot(kh, blk, co, N - 1) = hi;
          slot(kh, blk, co, N - 2) = lo;
        }
      }
  }
  p->layers = layers;
  if (G != T2_G && p->W >= T2_G) {
    p->layers_lat = layers;
    for (int L = pooled ? 1 : 0; L <= c.n_layers; ++L)
      if (!plan_segments(p->layers_lat[L], p->W, p->layers_lat[L].d, T2_G)) { p->layers_lat.clear(); break; }
  }
  HONK_TRY(up(p.get(), wpack, &p->d_wpack));
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
  if (pooled) {
    const int sm0 = S * (p->H0 + 2) * (p->W0 + 2) * static_cast<int>(sizeof(float));
    if (sm0 + 1024 > max_smem) return HONK_OK;
    HONK_TRY(set_smem(tc2_conv0_pool_mma_kernel<6>, sm0));
    if (p->conv0_tc) {
      HONK_TRY(set_smem(tc2_conv0_pool_tc_kernel<6, 3>, p->c0_smem));
      HONK_TRY(set_smem(tc2_conv0_pool_tc_kernel<4, 3>, p->c0_smem));
      HONK_TRY(set_smem(tc2_conv0_pool_tc_kernel<6, 2>, p->c0_smem));
      HONK_TRY(set_smem(tc2_conv0_pool_tc_kernel<4, 2>, p->c0_smem));
    }
    HONK_TRY(set_smem(tc2_conv0_pool_mma_kernel<4>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<6, 4, 3>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<4, 4, 3>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<6, 2, 2>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<4, 2, 2>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<6, 0, 0>, sm0));
    HONK_TRY(set_smem(tc2_conv0_pool_kernel<4, 0, 0>, sm0));
  }
  int per_sm = 0;
  if (ncg == 6) {
    HONK_TRY(set_smem(tc2_res_kernel<6, false>, p->smem_bytes));
    HONK_TRY(set_smem(tc2_res_kernel<6, true>, p->smem_bytes));
    HONK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tc2_res_kernel<6, false>, t2_threads(6), p->smem_bytes));
  } else {
    HONK_TRY(set_smem(tc2_res_kernel<4, false>, p->smem_bytes));
    HONK_TRY(set_smem(tc2_res_kernel<4, true>, p->smem_bytes));
    HONK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tc2_res_kernel<4, false>, t2_threads(4), p->smem_bytes));
  }
  p->max_resident = per_sm * m->n_sms;                 // CTAs that can be co-resident: the persistent grid never exceeds it
  if (p->max_resident < G) return HONK_OK;
  p->max_teams = std::min(p->max_teams, p->max_resident / G);
  *out = p.release();
  return HONK_OK;
}

// kernel "clips": groups of S stacked clips for the pooled models
static int64_t tc2_groups(const ResTc2Plan* p, int64_t n_clips) { return (n_clips + p->S - 1) / p->S; }
// Team size, groups in flight per team and segment table for a batch.  Throughput mode: the plan's (G, K).  When every
// group fits on its own team anyway, one group per team (K = 1) finishes sooner than pairs in lockstep; and a
// request-sized batch of a pooled model spreads each group over T2_G CTAs (latency mode: label(), /listen).
struct T2Mode {
  int G, K, max_teams;
  const std::vector<T2Layer>* layers;
};
static T2Mode tc2_mode(const ResTc2Plan* p, int64_t n_groups) {
  if (!p->tuned_by_env) {
    if (!p->layers_lat.empty() && n_groups <= p->n_sms / T2_G) return {T2_G, 1, p->n_sms / T2_G, &p->layers_lat};
    if (n_groups <= p->max_teams) return {p->G, 1, p->max_teams, &p->layers};
  }
  return {p->G, p->K, p->max_teams, &p->layers};
}
static int tc2_teams(const T2Mode& md, int64_t n_groups) {
  return static_cast<int>(std::min<int64_t>(md.max_teams, (n_groups + md.K - 1) / md.K));
}

int res_tc2_wave_clips(const ResTc2Plan* p) { return p->max_teams * p->K * p->S; }

int64_t res_tc2_workspace_bytes(const ResTc2Plan* p, int64_t n_clips) {
  const int64_t n_groups = tc2_groups(p, n_clips);
  const T2Mode md = tc2_mode(p, n_groups);
  const int n_slots = md.K * tc2_teams(md, n_groups);
  Arena a(nullptr, 0);
  a.take<unsigned char>(2ll * n_slots * p->slot_stride);
  a.take<float>(n_groups * p->S * p->NB);
  a.take<unsigned int>(n_groups);
  if (p->pool_h > 0) a.take<unsigned char>(n_groups * p->slot_stride);
  else a.take<unsigned char>(n_clips * p->x0_clip_stride);
  return align_up(a.off, 256);
}

unsigned char* res_tc2_x0_buffer(const ResTc2Plan* p, int64_t n_clips, void* ws, int64_t ws_bytes, int* H, int* W) {
  if (p->pool_h > 0 || !ws) return nullptr;
  const int64_t n_groups = tc2_groups(p, n_clips);
  const T2Mode md = tc2_mode(p, n_groups);
  const int n_slots = md.K * tc2_teams(md, n_groups);
  Arena ar(ws, ws_bytes);                               // the same sequence of takes as res_tc2_forward
  ar.take<unsigned char>(2ll * n_slots * p->slot_stride);
  ar.take<float>(n_groups * p->S * p->NB);
  ar.take<unsigned int>(n_groups);
  unsigned char* x0 = ar.take<unsigned char>(n_clips * p->x0_clip_stride);
  if (!ar.ok()) return nullptr;
  *H = p->H;
  *W = p->W;
  return x0;
}

int res_tc2_forward(const ResTc2Plan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                    cudaStream_t s) {
  const bool pooled = p->pool_h > 0;
  const int64_t n_groups = tc2_groups(p, n_clips);
  const T2Mode md = tc2_mode(p, n_groups);
  const int teams = tc2_teams(md, n_groups), n_slots = md.K * teams;
  Arena ar(ws, ws_bytes);
  unsigned char* act = ar.take<unsigned char>(2ll * n_slots * p->slot_stride);
  float* pool = ar.take<float>(n_groups * p->S * p->NB);
  unsigned int* flags = ar.take<unsigned int>(n_groups);
  const int64_t zero_bytes = ar.off;
  unsigned char* x0 = pooled ? ar.take<unsigned char>(n_groups * p->slot_stride) : ar.take<unsigned char>(n_clips * p->x0_clip_stride);
  HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: tcgen05 workspace overrun");
  {
    // What must start as zero: the T2_PAD rows above every channel plane (the convolutions' time padding), the separator rows
    // of stacked groups (never stored), the tail pad of every slot, the pool sums and the layer counters.  Real rows are
    // written by a layer before any later layer reads them, so the arena itself (66 MB for res15) is not cleared.
    ProfSpan span("res_zero", s);
    const long long planes_per_slot = static_cast<long long>(p->W) * p->ncg;
    const int tail_units = static_cast<int>((p->slot_stride - planes_per_slot * p->plane_rows * 16) / 16);
    const long long units = 2ll * n_slots * (planes_per_slot * (T2_PAD + p->S - 1) + tail_units);
    tc2_zero_pads_kernel<<<static_cast<unsigned>(std::min<long long>((units + 255) / 256, 16 * p->n_sms)), 256, 0, s>>>(
        act, p->slot_stride, 2 * n_slots, static_cast<int>(planes_per_slot), p->plane_rows, p->S, p->Hc, tail_units);
    HONK_LAUNCH_CHECK();
    HONK_CUDA(cudaMemsetAsync(pool, 0, static_cast<size_t>(zero_bytes - (reinterpret_cast<unsigned char*>(pool) - act)), s));
  }
  HONK_CHECK(d_mfcc || !pooled, HONK_ERR_INVALID, "internal: pooled models take MFCCs");
  if (pooled) {
    ProfSpan span("res_conv0_pool", s);
    const size_t smem = static_cast<size_t>(p->S) * (p->H0 + 2) * (p->W0 + 2) * sizeof(float);
    const dim3 grid(static_cast<unsigned>(n_groups));
#define T2_C0(NCG_, PH_, PW_)                                                                                                   \
  tc2_conv0_pool_kernel<NCG_, PH_, PW_><<<grid, NCG_ * 64, smem, s>>>(d_mfcc, p->d_w0, x0, static_cast<int>(n_clips), p->H0, p->W0, \
                                                                      p->pool_h, p->pool_w, p->Hc, p->W, p->S, p->plane_rows, p->slot_stride)
    const int pk = (p->pool_h == 4 && p->pool_w == 3) ? 1 : ((p->pool_h == 2 && p->pool_w == 2) ? 2 : 0);
    if (!p->conv0_f32 && p->conv0_tc) {
      const dim3 gtc(static_cast<unsigned>(std::min<long long>(n_groups, p->n_sms)));
      static const int c0_mode = getenv("HONK_TC2_C0_MODE") ? atoi(getenv("HONK_TC2_C0_MODE")) : 0;
#define T2_C0TC(NCG_, PW_)                                                                                                          \
  tc2_conv0_pool_tc_kernel<NCG_, PW_><<<gtc, T2C0_THREADS, p->c0_smem, s>>>(d_mfcc, reinterpret_cast<const unsigned char*>(p->d_w0t), x0, \
                                                                           static_cast<int>(n_clips), p->H0, p->pool_h, p->Hc, p->S,      \
                                                                           p->plane_rows, p->slot_stride, p->c0_img_bytes, c0_mode)
      if (p->ncg == 6) {
        if (p->pool_w == 3) T2_C0TC(6, 3); else T2_C0TC(6, 2);
      } else {
        if (p->pool_w == 3) T2_C0TC(4, 3); else T2_C0TC(4, 2);
      }
#undef T2_C0TC
    } else if (!p->conv0_f32) {
      if (p->ncg == 6)
        tc2_conv0_pool_mma_kernel<6><<<grid, 256, smem, s>>>(d_mfcc, p->d_w0b, x0, static_cast<int>(n_clips), p->H0, p->W0, p->pool_h,
                                                             p->pool_w, p->Hc, p->W, p->S, p->plane_rows, p->slot_stride);
      else
        tc2_conv0_pool_mma_kernel<4><<<grid, 256, smem, s>>>(d_mfcc, p->d_w0b, x0, static_cast<int>(n_clips), p->H0, p->W0, p->pool_h,
                                                             p->pool_w, p->Hc, p->W, p->S, p->plane_rows, p->slot_stride);
    } else if (p->ncg == 6) {
      if (pk == 1) T2_C0(6, 4, 3); else if (pk == 2) T2_C0(6, 2, 2); else T2_C0(6, 0, 0);
    } else {
      if (pk == 1) T2_C0(4, 4, 3); else if (pk == 2) T2_C0(4, 2, 2); else T2_C0(4, 0, 0);
    }
#undef T2_C0
    HONK_LAUNCH_CHECK();
  } else if (d_mfcc) {                                  // nullptr: the front end wrote x0 itself (fe_forward_x0)
    ProfSpan span("res_pack_x0", s);
    tc2_pack_x0_kernel<<<static_cast<unsigned>(n_clips), 256, p->H * (p->W + 1) * sizeof(float), s>>>(d_mfcc, x0, p->H, p->W);
    HONK_LAUNCH_CHECK();
  }
  T2Params a{};
  for (size_t i = 0; i < md.layers->size(); ++i) a.layers[i] = (*md.layers)[i];
  a.wpack = p->d_wpack;
  a.x0 = x0;
  a.y0 = x0;
  a.G = md.G;
  a.L0 = pooled ? 1 : 0;
  a.S = p->S;
  a.Hp1 = pooled ? p->Hc + 1 : p->H + 1;
  a.P = act;
  a.Q = act + static_cast<long long>(n_slots) * p->slot_stride;
  a.pool = pool;
  a.flags = flags;
  a.slot_stride = p->slot_stride;
  a.x0_clip_stride = p->x0_clip_stride;
  a.n_layers = p->n_layers;
  a.n_clips = static_cast<int>(n_groups);
  a.n_slots = n_slots;
  a.K = md.K;
  a.H = p->H;
  a.W = p->W;
  a.plane_rows = p->plane_rows;
  a.col_bytes = p->col_bytes;
  a.stage_stride = p->stage_stride;
  a.wbytes = p->wbytes;
  a.dbg = nullptr;
  a.mode = 0;
  static const int l2hint = getenv("HONK_TC2_L2HINT") ? atoi(getenv("HONK_TC2_L2HINT")) : 1;
  a.l2hint = l2hint;
  static long long* s_dbg = nullptr;
  const char* dbg_env = getenv("HONK_TC2_DEBUG");
  const bool dbg = dbg_env && atoi(dbg_env);
  if (dbg) {
    if (!s_dbg) HONK_CUDA(cudaMalloc(&s_dbg, 160 * T2_NDBG * sizeof(long long)));
    HONK_CUDA(cudaMemsetAsync(s_dbg, 0, 160 * T2_NDBG * sizeof(long long), s));
    a.dbg = s_dbg;
    if (const char* e = getenv("HONK_TC2_MODE")) a.mode = atoi(e);
  }
  // Experiment switch (HONK_TC2_L2PERSIST=1): mark the rotating activation buffers as persisting L2 lines for this launch.
  static const bool l2persist = [] { const char* e = getenv("HONK_TC2_L2PERSIST"); return e && atoi(e); }();
  if (l2persist) {
    static bool limit_set = false;
    if (!limit_set) {
      int dev = 0, max_persist = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
      cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, static_cast<size_t>(max_persist));
      fprintf(stderr, "tc2: persisting L2 limit %d MB\n", max_persist >> 20);
      limit_set = true;
    }
    int dev = 0, max_win = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    cudaStreamAttrValue av{};
    av.accessPolicyWindow.base_ptr = act;
    av.accessPolicyWindow.num_bytes = std::min<size_t>(static_cast<size_t>(2ll * n_slots * p->slot_stride), static_cast<size_t>(max_win));
    av.accessPolicyWindow.hitRatio = 1.0f;
    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    HONK_CUDA(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &av));
  }
  {
    ProfSpan span("res_conv", s);
    // The CTAs of a team spin on each other's layer counters, so the whole grid must be resident at once.  A cooperative
    // launch makes the driver schedule all CTAs together or not at all - also when another stream (a second serving
    // thread, the NCCL gather) holds SMs - instead of relying on an otherwise idle GPU; a grid that cannot be co-resident
    // is refused with an error rather than left to spin into the kernel's poll limit.
    HONK_CHECK(teams * md.G <= p->max_resident, HONK_ERR_UNSUPPORTED, "tc2_res_kernel: %d CTAs cannot be co-resident (limit %d)",
               teams * md.G, p->max_resident);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(teams * md.G);
    cfg.blockDim = dim3(t2_threads(p->ncg));
    cfg.dynamicSmemBytes = p->smem_bytes;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t le;
    if (p->ncg == 6) le = dbg ? cudaLaunchKernelEx(&cfg, tc2_res_kernel<6, true>, a) : cudaLaunchKernelEx(&cfg, tc2_res_kernel<6, false>, a);
    else le = dbg ? cudaLaunchKernelEx(&cfg, tc2_res_kernel<4, true>, a) : cudaLaunchKernelEx(&cfg, tc2_res_kernel<4, false>, a);
    HONK_CHECK(le == cudaSuccess, HONK_ERR_CUDA, "tc2_res_kernel cooperative launch failed: %s", cudaGetErrorString(le));
    HONK_LAUNCH_CHECK();
  }
  if (dbg) {
    std::vector<long long> h(160 * T2_NDBG);
    HONK_CUDA(cudaStreamSynchronize(s));
    HONK_CUDA(cudaMemcpy(h.data(), a.dbg, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
    double m[T2_NDBG] = {0};
    const int nb = teams * md.G;
    for (int b = 0; b < nb; ++b)
      for (int k = 0; k < T2_NDBG; ++k) m[k] += static_cast<double>(h[b * T2_NDBG + k]) / nb;
    fprintf(stderr, "tc2 dbg (mean cycles/CTA, %d CTAs, %lld clips): producer total %.0f flag-wait %.0f stage-empty-wait %.0f | "
            "mma total %.0f acc-empty-wait %.0f stage-full-wait %.0f weight-wait %.0f | epilogue total %.0f acc-full-wait %.0f | "
            "lookahead hits %.0f of %.0f tasks | f16 saturations %.0f\n",
            nb, static_cast<long long>(n_clips), m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[8], m[7], m[11], m[12], m[13] * nb);
  }
  if (l2persist) {
    cudaStreamAttrValue av{};
    av.accessPolicyWindow.num_bytes = 0;
    HONK_CUDA(cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &av));
  }
  {
    ProfSpan span("res_tail", s);
    const int total = static_cast<int>(n_clips) * p->n_labels;
    tc2_tail_kernel<<<(total + 127) / 128, 128, 0, s>>>(pool, p->d_wout, p->d_bout, d_logits, static_cast<int>(n_clips), p->NB,
                                                        p->n_labels, 1.0f / static_cast<float>(p->Hc * p->W));
    HONK_LAUNCH_CHECK();
  }
  return HONK_OK;
}

}  // namespace honk
