// FP32 CUDA-core operators: implicit-GEMM convolution / linear, pooling, BN + residual, mean,
// and the softmax/argmax label kernel.  See generic_ops.cuh.
#include "generic_ops.cuh"

#include <cstdlib>

namespace honk {

namespace {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;

// C[M,N] = A[M,K] * W[N,K]^T with A gathered on the fly from the NCHW input (im2col never
// materialised).  M = B*Ho*Wo, N = Cout, K = Cin*KH*KW.  64x64x16 tiles, 4x4 per thread.
__global__ void __launch_bounds__(256) conv_gemm_f32_kernel(const float* __restrict__ in,
                                                            const float* __restrict__ w,
                                                            const float* __restrict__ bias,
                                                            float* __restrict__ out, ConvGeom g, int relu) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const long long M = static_cast<long long>(g.B) * g.Ho * g.Wo;
  const int N = g.Cout, K = g.Cin * g.KH * g.KW, KHW = g.KH * g.KW;
  const int tid = threadIdx.x;
  const long long m0 = static_cast<long long>(blockIdx.x) * BM;
  const int n0 = blockIdx.y * BN;

  // A-load role: one m per thread, 4 k's
  const int la_m = tid % BM, la_k = tid / BM;           // la_k in 0..3
  const long long gm = m0 + la_m;
  const bool m_ok = gm < M;
  int hi0 = 0, wi0 = 0;
  const float* in_b = in;
  if (m_ok) {
    const int wo = static_cast<int>(gm % g.Wo);
    const long long r = gm / g.Wo;
    const int ho = static_cast<int>(r % g.Ho);
    const long long b = r / g.Ho;
    hi0 = ho * g.sh - g.ph;
    wi0 = wo * g.sw - g.pw;
    in_b = in + b * g.Cin * g.H * g.W;
  }
  // B-load role: k fastest
  const int lb_k = tid % BK, lb_n = tid / BK;            // lb_n in 0..15

  const int tm = (tid % 16) * TM, tn = (tid / 16) * TN;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int i = 0; i < BK / 4; ++i) {
      const int kk = la_k + 4 * i, k = k0 + kk;
      float v = 0.f;
      if (m_ok && k < K) {
        const int ci = k / KHW, rem = k - ci * KHW;
        const int kh = rem / g.KW, kw = rem - kh * g.KW;
        const int hi = hi0 + kh * g.dh, wi = wi0 + kw * g.dw;
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W)
          v = __ldg(in_b + (static_cast<long long>(ci) * g.H + hi) * g.W + wi);
      }
      As[kk][la_m] = v;
    }
#pragma unroll
    for (int i = 0; i < BN / 16; ++i) {
      const int nn = lb_n + 16 * i, n = n0 + nn, k = k0 + lb_k;
      Bs[lb_k][nn] = (n < N && k < K) ? __ldg(w + static_cast<long long>(n) * K + k) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[kk][tm + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[kk][tn + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  const long long HWo = static_cast<long long>(g.Ho) * g.Wo;
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const long long m = m0 + tm + i;
    if (m >= M) continue;
    const long long b = m / HWo, pos = m - b * HWo;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tn + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
      if (relu) v = fmaxf(v, 0.f);
      out[(b * N + n) * HWo + pos] = v;
    }
  }
}

// The same GEMM for large problems (training batches, the FP32 reference path at serving batch sizes): 128 x (16 TN) x 16
// tiles, 8 x TN outputs per thread, the next stage's global loads in flight under the current stage's FMAs (two shared-memory
// buffers, one barrier per stage).  Every output is still ONE fmaf chain over k in ascending order, so the result is
// bit-identical to conv_gemm_f32_kernel's.
constexpr int GB_BM = 128, GB_BK = 16, GB_MAXTAPS = 256;

template <int TN>
__global__ void __launch_bounds__(256) conv_gemm_f32_big_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                                const float* __restrict__ bias, float* __restrict__ out, ConvGeom g, int relu) {
  constexpr int BNT = 16 * TN;
  __shared__ __align__(16) float As[2][GB_BK][GB_BM + 4];
  __shared__ __align__(16) float Bs[2][GB_BK][BNT + 4];
  __shared__ short2 taps[GB_MAXTAPS];                          // (kh * dh, kw * dw) of tap r = kh * KW + kw
  const long long M = static_cast<long long>(g.B) * g.Ho * g.Wo;
  const int N = g.Cout, KHW = g.KH * g.KW, K = g.Cin * KHW;
  const int tid = threadIdx.x;
  const long long m0 = static_cast<long long>(blockIdx.x) * GB_BM;
  const int n0 = blockIdx.y * BNT;
  for (int r = tid; r < KHW; r += 256) taps[r] = make_short2(static_cast<short>((r / g.KW) * g.dh), static_cast<short>((r % g.KW) * g.dw));

  // A-load role: row la_m of the tile, the 8 consecutive k's [la_k0, la_k0 + 8) of every stage
  const int la_m = tid & 127, la_k0 = (tid >> 7) * 8;
  const long long gm = m0 + la_m;
  const bool m_ok = gm < M;
  int hi0 = 0, wi0 = 0;
  const float* in_b = in;
  if (m_ok) {
    const int wo = static_cast<int>(gm % g.Wo);
    const long long r = gm / g.Wo;
    const int ho = static_cast<int>(r % g.Ho);
    hi0 = ho * g.sh - g.ph;
    wi0 = wo * g.sw - g.pw;
    in_b = in + (r / g.Ho) * g.Cin * g.H * g.W;
  }
  int a_ci = la_k0 / KHW, a_r = la_k0 - a_ci * KHW;            // (channel, tap) of this thread's first k of the current stage
  // B-load role: k = tid % 16, rows tid / 16 + 16 j
  const int lb_k = tid & 15, lb_n = tid >> 4;
  // compute role: rows tx*4 + i and 64 + tx*4 + i, columns ty*TN + j
  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][TN];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float pa[8], pb[TN];
  auto fetch = [&](int k0) {
    int ci = a_ci, r = a_r;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v = 0.f;
      if (m_ok && k0 + la_k0 + i < K) {
        const short2 t = taps[r];
        const int hi = hi0 + t.x, wi = wi0 + t.y;
        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) v = __ldg(in_b + (static_cast<long long>(ci) * g.H + hi) * g.W + wi);
      }
      pa[i] = v;
      if (++r == KHW) { r = 0; ++ci; }
    }
    a_r += GB_BK;                                               // this thread's first k of the NEXT stage
    while (a_r >= KHW) { a_r -= KHW; ++a_ci; }
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + lb_n + 16 * j, k = k0 + lb_k;
      pb[j] = (n < N && k < K) ? __ldg(w + static_cast<long long>(n) * K + k) : 0.f;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 8; ++i) As[buf][la_k0 + i][la_m] = pa[i];
#pragma unroll
    for (int j = 0; j < TN; ++j) Bs[buf][lb_k][lb_n + 16 * j] = pb[j];
  };
  __syncthreads();                                              // taps
  fetch(0);
  stash(0);
  __syncthreads();
  const int n_stages = (K + GB_BK - 1) / GB_BK;
  for (int st = 0; st < n_stages; ++st) {
    const int buf = st & 1;
    if (st + 1 < n_stages) fetch((st + 1) * GB_BK);
#pragma unroll
    for (int kk = 0; kk < GB_BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][tx * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[TN];
#pragma unroll
      for (int j = 0; j < TN; ++j) b[j] = Bs[buf][kk][ty * TN + j];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (st + 1 < n_stages) stash(buf ^ 1);
    __syncthreads();
  }
  const long long HWo = static_cast<long long>(g.Ho) * g.Wo;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long m = m0 + (i < 4 ? tx * 4 + i : 64 + tx * 4 + (i - 4));
    if (m >= M) continue;
    const long long b = m / HWo, pos = m - b * HWo;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + ty * TN + j;
      if (n >= N) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
      if (relu) v = fmaxf(v, 0.f);
      out[(b * N + n) * HWo + pos] = v;
    }
  }
}

__global__ void pool2d_f32_kernel(const float* __restrict__ in, float* __restrict__ out, long long total, int H,
                                  int W, int Ho, int Wo, int kh, int kw, int is_max) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int wo = static_cast<int>(i % Wo);
  const long long r = i / Wo;
  const int ho = static_cast<int>(r % Ho);
  const long long bc = r / Ho;
  const float* p = in + (bc * H + static_cast<long long>(ho) * kh) * W + static_cast<long long>(wo) * kw;
  float acc = is_max ? -INFINITY : 0.f;
  for (int y = 0; y < kh; ++y)
    for (int x = 0; x < kw; ++x) {
      const float v = __ldg(p + y * W + x);
      acc = is_max ? fmaxf(acc, v) : acc + v;
    }
  out[i] = is_max ? acc : acc / static_cast<float>(kh * kw);
}

// (res and res_out are the same buffer when the residual stream is updated in place: no __restrict__ on them)
__global__ void bn_residual_f32_kernel(const float* __restrict__ y, const float* res,
                                       float* res_out, float* __restrict__ out,
                                       const float* __restrict__ mean, const float* __restrict__ scale,
                                       long long total, int C, int HW) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = static_cast<int>((i / HW) % C);
  float x = y[i];
  if (res) x += res[i];
  if (res_out) res_out[i] = x;
  out[i] = (x - __ldg(mean + c)) * __ldg(scale + c);
}

__global__ void mean_hw_f32_kernel(const float* __restrict__ in, float* __restrict__ out, int HW) {
  const long long bc = blockIdx.x;
  const float* p = in + bc * HW;
  float s = 0.f;
  for (int i = threadIdx.x; i < HW; i += blockDim.x) s += p[i];
  __shared__ float red[32];
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) out[bc] = s / static_cast<float>(HW);
  }
}

// softmax -> (argmax, max prob); one warp per clip (service.py:103-104)
__global__ void label_kernel(const float* __restrict__ logits, long long n, int L, int* __restrict__ labels,
                             float* __restrict__ probs) {
  const long long clip = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (clip >= n) return;
  const float* z = logits + clip * L;
  float best = -INFINITY;
  int arg = 0x7fffffff;
  bool nan = false;
  for (int i = lane; i < L; i += 32) {
    const float v = z[i];
    nan |= (v != v);
    if (v > best || arg == 0x7fffffff) { best = v; arg = i; }       // a row of -inf still names its first entry
  }
  for (int o = 16; o; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
  }
  nan = __any_sync(0xffffffffu, nan);
  float sum = 0.f;
  for (int i = lane; i < L; i += 32) sum += expf(z[i] - best);
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) {
    // one NaN logit makes every softmax entry NaN: np.argmax then names entry 0 and max() is NaN (service.py:103-104)
    if (labels) labels[clip] = nan ? 0 : arg;
    if (probs) probs[clip] = nan ? __int_as_float(0x7fc00000) : 1.f / sum;
  }
}

// print_eval's argmax == label count and CrossEntropyLoss's per-clip term (utils/train.py:41-46, 66-83): one warp
// per clip, block-level reduction in a fixed order, one atomic per block.
__global__ void __launch_bounds__(256) eval_metrics_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                           long long n, int L, int* __restrict__ pred,
                                                           unsigned long long* __restrict__ correct, double* __restrict__ loss_sum) {
  __shared__ double s_loss[8];
  __shared__ int s_ok[8];
  const long long clip = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  double loss = 0.0;
  int ok = 0;
  if (clip < n) {
    const float* z = logits + clip * L;
    float best = -INFINITY;
    int arg = 0x7fffffff, first_nan = 0x7fffffff;
    for (int i = lane; i < L; i += 32) {
      const float v = z[i];
      if (v != v && first_nan == 0x7fffffff) first_nan = i;
      if (v > best || arg == 0x7fffffff) { best = v; arg = i; }
    }
    for (int o = 16; o; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }      // first maximum, like torch.max
      first_nan = min(first_nan, __shfl_xor_sync(0xffffffffu, first_nan, o));
    }
    if (first_nan != 0x7fffffff) { arg = first_nan; best = __int_as_float(0x7fc00000); }   // torch.max propagates NaN: its index wins
    double sum = 0.0;
    for (int i = lane; i < L; i += 32) sum += exp(static_cast<double>(z[i]) - static_cast<double>(best));
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const long long y = labels[clip];
    if (lane == 0) {
      if (pred) pred[clip] = arg;
      ok = (static_cast<long long>(arg) == y);
      // logsumexp(z) - z[y]; a target outside [0, L) is an error in torch (device assert): it poisons the sum with NaN here
      // instead of being skipped silently, and honk_b200.train.evaluate raises on it
      loss = (y >= 0 && y < L) ? log(sum) + static_cast<double>(best) - static_cast<double>(z[y]) : __longlong_as_double(0x7ff8000000000000LL);
    }
  }
  if (lane == 0) { s_loss[wib] = loss; s_ok[wib] = ok; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    int c = 0;
    for (int k = 0; k < 8; ++k) { t += s_loss[k]; c += s_ok[k]; }
    atomicAdd(loss_sum, t);
    atomicAdd(correct, static_cast<unsigned long long>(c));
  }
}

// SpeechDataset.load_audio's arithmetic (utils/model.py:281-303) for a batch, given the random draws:
// zero-pad to in_len (:297), time shift (:267-273: out[t] = x[t + shift], zero outside), then, where gain >= 0,
// clip(gain * bg[off + t] + x, -1, 1) in float32 with separately rounded multiply and add like numpy (:299-303).
__global__ void __launch_bounds__(256) load_audio_kernel(const float* __restrict__ x, long long x_stride, const int* __restrict__ lens,
                                                         const int* __restrict__ shifts, const float* __restrict__ gains,
                                                         const long long* __restrict__ noise_off, const float* __restrict__ bg,
                                                         float* __restrict__ out, int in_len) {
  const long long clip = blockIdx.y;
  const int len = lens ? min(lens[clip], in_len) : in_len;        // samples past in_len never reach the output
  const int shift = shifts ? shifts[clip] : 0;
  const float g = gains ? gains[clip] : -1.f;
  const bool noisy = g >= 0.f;                                    // NaN or negative: no noise, no clip
  const float* xi = x + clip * x_stride;
  const float* bi = (noisy && bg) ? bg + noise_off[clip] : nullptr;
  float* o = out + clip * in_len;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < in_len; t += gridDim.x * blockDim.x) {
    const int src = t + shift;
    float v = (src >= 0 && src < len) ? __ldg(xi + src) : 0.f;
    if (noisy) {
      const float nz = bi ? __fmul_rn(g, __ldg(bi + t)) : 0.f;
      v = fminf(fmaxf(__fadd_rn(nz, v), -1.f), 1.f);
    }
    o[t] = v;
  }
}

}  // namespace

int conv2d_f32(const float* in, const float* w, const float* bias, float* out, const ConvGeom& g, bool relu,
               cudaStream_t s) {
  const long long M = static_cast<long long>(g.B) * g.Ho * g.Wo;
  if (M == 0) return HONK_OK;
  HONK_CHECK(g.Ho > 0 && g.Wo > 0, HONK_ERR_INVALID, "convolution output is empty");
  static const bool no_big = getenv("HONK_F32_SMALL_TILES") && atoi(getenv("HONK_F32_SMALL_TILES")) != 0;
  if (!no_big && M >= 8192 && g.Cout >= 16 && g.Cin * g.KH * g.KW >= 16 && g.KH * g.KW <= GB_MAXTAPS && g.KH * g.dh < 32768 &&
      g.KW * g.dw < 32768) {
    const int tn = g.Cout > 48 ? 4 : (g.Cout > 32 ? 3 : 2);
    dim3 gb(static_cast<unsigned>((M + GB_BM - 1) / GB_BM), ceil_div(g.Cout, 16 * tn));
    if (tn == 4) conv_gemm_f32_big_kernel<4><<<gb, 256, 0, s>>>(in, w, bias, out, g, relu ? 1 : 0);
    else if (tn == 3) conv_gemm_f32_big_kernel<3><<<gb, 256, 0, s>>>(in, w, bias, out, g, relu ? 1 : 0);
    else conv_gemm_f32_big_kernel<2><<<gb, 256, 0, s>>>(in, w, bias, out, g, relu ? 1 : 0);
    HONK_LAUNCH_CHECK();
    return HONK_OK;
  }
  dim3 grid(static_cast<unsigned>((M + BM - 1) / BM), ceil_div(g.Cout, BN));
  conv_gemm_f32_kernel<<<grid, 256, 0, s>>>(in, w, bias, out, g, relu ? 1 : 0);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

int pool2d_f32(const float* in, float* out, int BC, int H, int W, int kh, int kw, bool is_max, cudaStream_t s) {
  const int Ho = H / kh, Wo = W / kw;
  const long long total = static_cast<long long>(BC) * Ho * Wo;
  if (total == 0) return HONK_OK;
  pool2d_f32_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(in, out, total, H, W, Ho, Wo, kh,
                                                                              kw, is_max ? 1 : 0);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

int bn_residual_f32(const float* y, const float* res, float* res_out, float* out, const float* mean,
                    const float* scale, int B, int C, int HW, cudaStream_t s) {
  const long long total = static_cast<long long>(B) * C * HW;
  if (total == 0) return HONK_OK;
  bn_residual_f32_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(y, res, res_out, out, mean,
                                                                                   scale, total, C, HW);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

int mean_hw_f32(const float* in, float* out, int BC, int HW, cudaStream_t s) {
  if (BC == 0) return HONK_OK;
  mean_hw_f32_kernel<<<BC, 256, 0, s>>>(in, out, HW);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

}  // namespace honk

using namespace honk;

extern "C" int honk_label_from_logits(const float* d_logits, int64_t n_clips, int32_t n_labels, int32_t* d_labels,
                                      float* d_probs, void* stream) {
  HONK_CHECK(n_clips >= 0 && n_labels >= 1, HONK_ERR_INVALID, "bad n_clips/n_labels");
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(d_logits, HONK_ERR_INVALID, "null logits");
  const long long threads = n_clips * 32;
  label_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_logits, n_clips, n_labels, d_labels, d_probs);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

extern "C" int honk_eval_metrics(const float* d_logits, const int64_t* d_labels, int64_t n_clips, int32_t n_labels,
                                 int32_t* d_pred, int64_t* d_correct, double* d_loss_sum, void* stream) {
  HONK_CHECK(n_clips >= 0 && n_labels >= 1, HONK_ERR_INVALID, "bad n_clips/n_labels");
  HONK_CHECK(d_correct && d_loss_sum, HONK_ERR_INVALID, "honk_eval_metrics: null output");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  HONK_CUDA(cudaMemsetAsync(d_correct, 0, sizeof(int64_t), s));
  HONK_CUDA(cudaMemsetAsync(d_loss_sum, 0, sizeof(double), s));
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(d_logits && d_labels, HONK_ERR_INVALID, "honk_eval_metrics: null input");
  eval_metrics_kernel<<<static_cast<unsigned>((n_clips + 7) / 8), 256, 0, s>>>(
      d_logits, reinterpret_cast<const long long*>(d_labels), n_clips, n_labels, d_pred,
      reinterpret_cast<unsigned long long*>(d_correct), d_loss_sum);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

extern "C" int honk_load_audio(const float* d_pcm, int64_t n_clips, int64_t clip_stride, const int32_t* d_lengths,
                               const int32_t* d_shifts, const float* d_noise_gain, const int64_t* d_noise_offset,
                               const float* d_bg_noise, int32_t input_length, float* d_out, void* stream) {
  HONK_CHECK(n_clips >= 0 && n_clips <= 65535 && input_length >= 1, HONK_ERR_INVALID, "bad n_clips (<= 65535) / input_length");
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(d_pcm && d_out, HONK_ERR_INVALID, "honk_load_audio: null buffer");
  HONK_CHECK(!d_noise_gain || !d_bg_noise || d_noise_offset, HONK_ERR_INVALID, "noise offsets missing");
  dim3 grid(static_cast<unsigned>(std::min(16, (input_length + 1023) / 1024)), static_cast<unsigned>(n_clips));
  load_audio_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_pcm, clip_stride, d_lengths, d_shifts, d_noise_gain, reinterpret_cast<const long long*>(d_noise_offset), d_bg_noise,
      d_out, input_length);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}
