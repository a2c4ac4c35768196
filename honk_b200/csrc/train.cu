// One SGD training step on the GPU (reference: utils/train.py:127-139 - model.train(), forward, CrossEntropyLoss, backward,
// torch.optim.SGD.step) over SpeechResModel (utils/model.py:82-121) and SpeechModel (utils/model.py:123-205; second half of
// this file: conv + bias + ReLU -> Dropout -> MaxPool, Linear layers with Dropout).
//
// All arithmetic is FP32 on the CUDA cores (reductions accumulate in FP64) so that the step agrees with the reference's
// FP32 autograd to summation-order noise; nothing here calls a library.  The step is, per layer i = 1..n,
//   y_i = relu(conv_i(x_{i-1}));  r_i = y_i (+ old, i even; old = r_i);  x_i = BatchNorm_train(r_i)     (model.py:107-117)
// with BatchNorm2d(affine=False) in TRAINING mode: batch mean / biased variance normalise, and the running statistics
// move by momentum 0.1 towards the batch mean / UNBIASED variance (torch.nn.BatchNorm2d defaults, model.py:98).
// Backward keeps x_i and y_i of every layer (2 x [B,C,H,W] floats per layer), so the workspace is sized per batch.
#include "model.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

using namespace honk;

namespace {

constexpr float kBnEps = 1e-5f, kBnMomentum = 0.1f;
constexpr int kStatSplit = 64;   // partial sums per channel in the BatchNorm reductions

// ---------------------------------------------------------------------------------------------------------------------
// elementwise / reduction kernels
// ---------------------------------------------------------------------------------------------------------------------
__global__ void add_f32_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = a[i] + b[i];
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double t = 0.0;
  for (int i = 0; i < (blockDim.x >> 5); ++i) t += sh[i];
  return t;
}

// grid (C, kStatSplit): partial[c][split] = {sum a, sum a*b (b == nullptr: a*a)} over the split's share of (clip, pos).
__global__ void __launch_bounds__(256) chan_sums_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                        double* __restrict__ partial, int B, int C, int HW) {
  __shared__ double sh[8];
  const int c = blockIdx.x, sp = blockIdx.y;
  const long long total = static_cast<long long>(B) * HW;
  const long long per = (total + gridDim.y - 1) / gridDim.y;
  const long long lo = sp * per, hi = min(total, lo + per);
  double s1 = 0.0, s2 = 0.0;
  for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const long long bb = i / HW, pos = i - bb * HW;
    const long long idx = (bb * C + c) * HW + pos;
    const float va = a[idx];
    const float vb = b ? b[idx] : va;
    s1 += va;
    s2 += static_cast<double>(va) * vb;
  }
  s1 = block_sum(s1, sh);
  s2 = block_sum(s2, sh);
  if (threadIdx.x == 0) {
    partial[(static_cast<long long>(c) * gridDim.y + sp) * 2 + 0] = s1;
    partial[(static_cast<long long>(c) * gridDim.y + sp) * 2 + 1] = s2;
  }
}

// One thread per channel: batch mean, biased variance -> mean[c], invstd[c]; running statistics updated in place.
__global__ void bn_train_finalize_kernel(const double* __restrict__ partial, int nsplit, long long count, float* __restrict__ mean,
                                         float* __restrict__ invstd, float* __restrict__ running_mean,
                                         float* __restrict__ running_var, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s1 = 0.0, s2 = 0.0;
  for (int i = 0; i < nsplit; ++i) {
    s1 += partial[(static_cast<long long>(c) * nsplit + i) * 2 + 0];
    s2 += partial[(static_cast<long long>(c) * nsplit + i) * 2 + 1];
  }
  const double m = s1 / count;
  double var = s2 / count - m * m;
  if (var < 0.0) var = 0.0;
  mean[c] = static_cast<float>(m);
  invstd[c] = static_cast<float>(1.0 / sqrt(var + static_cast<double>(kBnEps)));
  const double unbiased = count > 1 ? var * (static_cast<double>(count) / (count - 1)) : var;
  running_mean[c] = (1.f - kBnMomentum) * running_mean[c] + kBnMomentum * static_cast<float>(m);
  running_var[c] = (1.f - kBnMomentum) * running_var[c] + kBnMomentum * static_cast<float>(unbiased);
}

__global__ void bn_apply_kernel(const float* __restrict__ r, const float* __restrict__ mean, const float* __restrict__ invstd,
                                float* __restrict__ out, int C, int HW, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i / HW) % C);
    out[i] = (r[i] - mean[c]) * invstd[c];
  }
}

// sums[c] = {mean_b(dx), mean_b(dx * xhat)} from the partials of chan_sums_kernel(dx, xhat)
__global__ void bn_bwd_finalize_kernel(const double* __restrict__ partial, int nsplit, long long count, float* __restrict__ sums, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s1 = 0.0, s2 = 0.0;
  for (int i = 0; i < nsplit; ++i) {
    s1 += partial[(static_cast<long long>(c) * nsplit + i) * 2 + 0];
    s2 += partial[(static_cast<long long>(c) * nsplit + i) * 2 + 1];
  }
  sums[2 * c + 0] = static_cast<float>(s1 / count);
  sums[2 * c + 1] = static_cast<float>(s2 / count);
}

// dr = invstd * (dx - mean(dx) - xhat * mean(dx * xhat))  [+ dold_in];  dold_out = dr (residual layers);
// dc = dr * (y > 0)  (gradient at the convolution's output, before the ReLU).
__global__ void bn_relu_bwd_kernel(const float* __restrict__ dx, const float* __restrict__ xhat, const float* __restrict__ y,
                                   const float* __restrict__ sums, const float* __restrict__ invstd,
                                   const float* dold_in, float* dold_out /* may be the same buffer */, float* __restrict__ dc,
                                   int C, int HW, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>((i / HW) % C);
    float dr = invstd[c] * (dx[i] - sums[2 * c] - xhat[i] * sums[2 * c + 1]);
    if (dold_in) dr += dold_in[i];
    if (dold_out) dold_out[i] = dr;
    dc[i] = y[i] > 0.f ? dr : 0.f;
  }
}

// conv0's side: g = da (+ dold); through the optional AvgPool2d (floor mode) and the ReLU of y0 [B*C, H, W]
__global__ void stem_bwd_kernel(const float* __restrict__ da, const float* __restrict__ dold, const float* __restrict__ y0,
                                float* __restrict__ dc0, int H, int W, int ph, int pw, int Hp, int Wp, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int w = static_cast<int>(i % W);
    const int h = static_cast<int>((i / W) % H);
    const long long bc = i / (static_cast<long long>(W) * H);
    float g = 0.f;
    const int hp = h / ph, wp = w / pw;
    if (hp < Hp && wp < Wp) {
      const long long j = (bc * Hp + hp) * Wp + wp;
      g = da[j] + (dold ? dold[j] : 0.f);
      g /= static_cast<float>(ph * pw);
    }
    dc0[i] = y0[i] > 0.f ? g : 0.f;
  }
}

// wt[ci][co][KH-1-kh][KW-1-kw] = w[co][ci][kh][kw]: the data gradient of a stride-1 convolution is the same kind of
// convolution of dy with these weights (pad' = dilation * (K - 1) - pad, which equals pad for the res* layers).
__global__ void flip_weights_kernel(const float* __restrict__ w, float* __restrict__ wt, int Cout, int Cin, int KHW) {
  const int n = Cout * Cin * KHW;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int tap = i % KHW, ci = (i / KHW) % Cin, co = i / (KHW * Cin);
    wt[(ci * Cout + co) * KHW + (KHW - 1 - tap)] = w[i];
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// weight gradient of a convolution (any kernel size, stride, padding, dilation) as a split-K FP32 GEMM:
//   dw[co][(ci, kh, kw)] = sum_pos dy[b][co][ho][wo] * x[b][ci][ho*sh - ph + kh*dh][wo*sw - pw + kw*dw],   pos = (b, ho, wo)
// M = Cout (tile BM), N = Cin*KH*KW (tile 64), K = B*Ho*Wo output positions, BK = 16 per stage.  grid (N tiles, splits,
// M tiles); each block reduces its contiguous share of the positions; partials go to part[split][co][n] and
// wgrad_reduce_kernel adds them in a fixed order (deterministic).  Both operands are read with the position as the
// fastest index across a half-warp, i.e. 64-byte segments along w.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int WG_BN = 64, WG_BK = 16, WG_THREADS = 256;

template <int BM>
__global__ void __launch_bounds__(WG_THREADS) wgrad_gemm_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                float* __restrict__ part, ConvGeom g, long long k_per_split) {
  constexpr int TM = BM / 16, TN = 4, LA = BM / 16;            // LA: A loads per thread per stage
  __shared__ float As[WG_BK][BM + 4];
  __shared__ float Bs[WG_BK][WG_BN + 4];
  const int tid = threadIdx.x;
  const int HWo = g.Ho * g.Wo, HWi = g.H * g.W, KHW = g.KH * g.KW, N = g.Cin * KHW, Cout = g.Cout;
  const long long K = static_cast<long long>(g.B) * HWo;
  const int n0 = blockIdx.x * WG_BN, m0 = blockIdx.z * BM, sp = blockIdx.y;
  const long long k_lo = sp * k_per_split, k_hi = min(K, k_lo + k_per_split);
  // load roles: position kk = tid % 16 of the stage; rows tid / 16 + 16 * i of the tile
  const int kk = tid & 15, row = tid >> 4;
  // this thread's output position (b, ho, wo), advanced by 16 per stage
  long long pos = k_lo + kk;
  int b = static_cast<int>(pos / HWo);
  int rem = static_cast<int>(pos - static_cast<long long>(b) * HWo);
  int ho = rem / g.Wo, wo = rem - ho * g.Wo;
  // B-operand columns of this thread: (ci, row offset, column offset) for n = n0 + row + 16 * i
  int b_ci[4], b_dh[4], b_dw[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int n = n0 + row + 16 * i;
    if (n < N) {
      const int ci = n / KHW, tap = n - ci * KHW, kh = tap / g.KW, kw = tap - kh * g.KW;
      b_ci[i] = ci;
      b_dh[i] = kh * g.dh - g.ph;
      b_dw[i] = kw * g.dw - g.pw;
    } else {
      b_ci[i] = -1; b_dh[i] = 0; b_dw[i] = 0;
    }
  }
  const int tm = (tid & 15) * TM, tn = (tid >> 4) * TN;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  for (long long k0 = k_lo; k0 < k_hi; k0 += WG_BK) {
    const bool k_ok = k0 + kk < k_hi;
    float av[LA], bv[4];
#pragma unroll
    for (int i = 0; i < LA; ++i) {
      const int m = m0 + row + 16 * i;
      av[i] = (k_ok && m < Cout) ? __ldg(dy + (static_cast<long long>(b) * Cout + m) * HWo + ho * g.Wo + wo) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int hh = ho * g.sh + b_dh[i], ww = wo * g.sw + b_dw[i];
      bv[i] = (k_ok && b_ci[i] >= 0 && hh >= 0 && hh < g.H && ww >= 0 && ww < g.W)
                  ? __ldg(x + (static_cast<long long>(b) * g.Cin + b_ci[i]) * HWi + hh * g.W + ww) : 0.f;
    }
    __syncthreads();                       // the previous stage's reads are done
#pragma unroll
    for (int i = 0; i < LA; ++i) As[kk][row + 16 * i] = av[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) Bs[kk][row + 16 * i] = bv[i];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < WG_BK; ++q) {
      float a[TM], bb[TN];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[q][tm + i];
#pragma unroll
      for (int j = 0; j < TN; ++j) bb[j] = Bs[q][tn + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    // advance this thread's position by one stage
    wo += WG_BK;
    while (wo >= g.Wo) { wo -= g.Wo; ++ho; }
    while (ho >= g.Ho) { ho -= g.Ho; ++b; }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + tm + i;
    if (m >= Cout) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int n = n0 + tn + j;
      if (n < N) part[(static_cast<long long>(sp) * Cout + m) * N + n] = acc[i][j];
    }
  }
}

// The same GEMM with (16 TM) x 128 x 16 tiles, TM x 8 outputs per thread and the next stage's loads in flight under the FMAs
// (two shared-memory buffers, one barrier per stage): used when there are at least 128 (ci, tap) columns.  Each partial sum is
// still one fmaf chain over the split's positions in ascending order (the number of splits may differ from the small kernel's).
constexpr int WGB_BN = 128;

template <int TM>
__global__ void __launch_bounds__(WG_THREADS) wgrad_gemm_big_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                    float* __restrict__ part, ConvGeom g, long long k_per_split) {
  constexpr int BM = 16 * TM;
  __shared__ __align__(16) float As[2][WG_BK][BM + 4];
  __shared__ __align__(16) float Bs[2][WG_BK][WGB_BN + 4];
  const int tid = threadIdx.x;
  const int HWo = g.Ho * g.Wo, HWi = g.H * g.W, KHW = g.KH * g.KW, N = g.Cin * KHW, Cout = g.Cout;
  const long long K = static_cast<long long>(g.B) * HWo;
  const int n0 = blockIdx.x * WGB_BN, m0 = blockIdx.z * BM, sp = blockIdx.y;
  const long long k_lo = sp * k_per_split, k_hi = min(K, k_lo + k_per_split);
  const int kk = tid & 15, row = tid >> 4;                       // load roles: position kk of the stage, rows / columns row + 16 i
  long long pos = k_lo + kk;
  int b = static_cast<int>(pos / HWo);
  int rem = static_cast<int>(pos - static_cast<long long>(b) * HWo);
  int ho = rem / g.Wo, wo = rem - ho * g.Wo;
  int b_off[8], b_d[8];                                          // per column: ci * H * W, and (row offset << 16 | column offset + 32768)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int n = n0 + row + 16 * i;
    if (n < N) {
      const int ci = n / KHW, tap = n - ci * KHW, kh = tap / g.KW, kw = tap - kh * g.KW;
      b_off[i] = ci * HWi;
      b_d[i] = ((kh * g.dh - g.ph + 16384) << 16) | (kw * g.dw - g.pw + 16384);
    } else {
      b_off[i] = -1; b_d[i] = 0;
    }
  }
  const int tx = tid & 15, ty = tid >> 4;                        // compute roles: columns tx*4 + j and 64 + tx*4 + j, rows ty*TM + i
  float acc[TM][8];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  float pa[TM], pb[8];
  auto fetch = [&](long long k0) {
    const bool k_ok = k0 + kk < k_hi;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int m = m0 + row + 16 * i;
      pa[i] = (k_ok && m < Cout) ? __ldg(dy + (static_cast<long long>(b) * Cout + m) * HWo + ho * g.Wo + wo) : 0.f;
    }
    const float* xb = x + static_cast<long long>(b) * g.Cin * HWi;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int hh = ho * g.sh + (b_d[i] >> 16) - 16384, ww = wo * g.sw + (b_d[i] & 0xFFFF) - 16384;
      pb[i] = (k_ok && b_off[i] >= 0 && hh >= 0 && hh < g.H && ww >= 0 && ww < g.W) ? __ldg(xb + b_off[i] + hh * g.W + ww) : 0.f;
    }
    wo += WG_BK;                                                 // this thread's position of the NEXT stage
    while (wo >= g.Wo) { wo -= g.Wo; ++ho; }
    while (ho >= g.Ho) { ho -= g.Ho; ++b; }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < TM; ++i) As[buf][kk][row + 16 * i] = pa[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) Bs[buf][kk][row + 16 * i] = pb[i];
  };
  fetch(k_lo);
  stash(0);
  __syncthreads();
  int st = 0;
  for (long long k0 = k_lo; k0 < k_hi; k0 += WG_BK, ++st) {
    const int buf = st & 1;
    const bool more = k0 + WG_BK < k_hi;
    if (more) fetch(k0 + WG_BK);
#pragma unroll
    for (int q = 0; q < WG_BK; ++q) {
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][q][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][q][64 + tx * 4]);
      const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float a[TM];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[buf][q][ty * TM + i];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    if (more) stash(buf ^ 1);
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= Cout) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n < N) part[(static_cast<long long>(sp) * Cout + m) * N + n] = acc[i][j];
    }
  }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ dw, int nsplit, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int k = 0; k < nsplit; ++k) s += part[static_cast<long long>(k) * n + i];
  dw[i] = static_cast<float>(s);
}

// ---------------------------------------------------------------------------------------------------------------------
// head: CrossEntropyLoss (mean) forward + gradient, Linear backward, mean-pool backward
// ---------------------------------------------------------------------------------------------------------------------
// One thread per clip: dlogits = (softmax - onehot) / B; rowloss[b] = logsumexp - logit[target] (NaN for a bad target).
__global__ void ce_fwd_bwd_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, float* __restrict__ dlogits,
                                  float* __restrict__ rowloss, int B, int L) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float* z = logits + static_cast<long long>(b) * L;
  float mx = z[0];
  for (int l = 1; l < L; ++l) mx = fmaxf(mx, z[l]);
  float se = 0.f;
  for (int l = 0; l < L; ++l) se += expf(z[l] - mx);
  const float lse = mx + logf(se);
  const long long t = labels[b];
  const bool ok = t >= 0 && t < L;
  rowloss[b] = ok ? lse - z[t] : __int_as_float(0x7fc00000);
  const float inv = 1.f / static_cast<float>(B);
  for (int l = 0; l < L; ++l) dlogits[static_cast<long long>(b) * L + l] = (expf(z[l] - lse) - ((ok && l == t) ? 1.f : 0.f)) * inv;
}

// One block: loss = mean(rowloss) in FP64, summed in index order.
__global__ void loss_mean_kernel(const float* __restrict__ rowloss, float* __restrict__ loss, int B) {
  __shared__ double sh[8];
  double s = 0.0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) s += rowloss[i];
  s = block_sum(s, sh);
  if (threadIdx.x == 0) *loss = static_cast<float>(s / B);
}

// dW[l][c] = sum_b dlogits[b][l] * pooled[b][c];  db[l] = sum_b dlogits[b][l]   (one thread per (l, c); c == C is the bias)
__global__ void linear_wgrad_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled, float* __restrict__ dW,
                                    float* __restrict__ db, int B, int L, int C) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L * (C + 1)) return;
  const int l = i / (C + 1), c = i - l * (C + 1);
  double s = 0.0;
  for (int b = 0; b < B; ++b) s += static_cast<double>(dlogits[static_cast<long long>(b) * L + l]) * (c < C ? pooled[static_cast<long long>(b) * C + c] : 1.f);
  if (c < C) dW[l * C + c] = static_cast<float>(s);
  else db[l] = static_cast<float>(s);
}

// dx[b][c][:] = (sum_l dlogits[b][l] * W[l][c]) / HW     (gradient of mean over the map followed by the Linear layer)
__global__ void head_dgrad_kernel(const float* __restrict__ dlogits, const float* __restrict__ Wout, float* __restrict__ dx, int L, int C, int HW) {
  const long long bc = blockIdx.x;
  const int c = static_cast<int>(bc % C);
  const long long b = bc / C;
  float s = 0.f;
  for (int l = 0; l < L; ++l) s = fmaf(dlogits[b * L + l], Wout[l * C + c], s);
  s /= static_cast<float>(HW);
  for (int i = threadIdx.x; i < HW; i += blockDim.x) dx[bc * HW + i] = s;
}

// torch.optim.SGD.step (weight_decay, momentum with dampening 0, optional Nesterov); first == 1: momentum buffer = grad.
__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ mom, int n, float lr, float momentum,
                           float weight_decay, int nesterov, int first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float grad = g[i];
  if (weight_decay != 0.f) grad = fmaf(weight_decay, p[i], grad);
  if (momentum != 0.f) {
    const float buf = first ? grad : fmaf(momentum, mom[i], grad);
    mom[i] = buf;
    grad = nesterov ? fmaf(momentum, buf, grad) : buf;
  }
  p[i] = p[i] - lr * grad;
}

// ---------------------------------------------------------------------------------------------------------------------
// cnn-* family (SpeechModel, utils/model.py:123-205): Dropout, MaxPool with argmax, Linear layers
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mix32(uint32_t x) {           // lowbias32 integer hash
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

// nn.Dropout in training mode: y = x * keep / (1 - p).  keep comes from the caller's mask (uint8, what the parity tests use) or
// is drawn here (counter-based hash of seed, site and element; the drawn mask is written to `mask` for the backward pass).
// (x and y may be the same buffer: no __restrict__ on them)
__global__ void dropout_fwd_kernel(const float* x, const unsigned char* __restrict__ user_mask, unsigned char* __restrict__ mask,
                                   float* y, long long n, float p, float scale, uint32_t seed) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    unsigned char k;
    if (user_mask) k = user_mask[i] != 0;
    else k = (mix32(seed ^ mix32(static_cast<uint32_t>(i) + 0x9e3779b9u * static_cast<uint32_t>(i >> 32))) >> 8) * (1.f / 16777216.f) >= p;
    mask[i] = k;
    y[i] = k ? x[i] * scale : 0.f;
  }
}

// dx = dy * keep / (1 - p), then through the ReLU that produced the dropped tensor when `relu_out` (the ReLU's output) is given
// (dy and dx may be the same buffer: no __restrict__ on them)
__global__ void dropout_relu_bwd_kernel(const float* dy, const unsigned char* __restrict__ mask, const float* __restrict__ relu_out,
                                        float* dx, long long n, float scale) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float g = (!mask || mask[i]) ? dy[i] * scale : 0.f;
    if (relu_out && !(relu_out[i] > 0.f)) g = 0.f;
    dx[i] = g;
  }
}

// nn.MaxPool2d(kernel = stride, floor mode) with the flat index (within the plane) of the FIRST maximum of each window in
// scan order, which is the element torch's backward sends the gradient to.
__global__ void maxpool_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, int* __restrict__ idx, long long total, int H, int W,
                                   int Ho, int Wo, int kh, int kw) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int wo = static_cast<int>(i % Wo), ho = static_cast<int>((i / Wo) % Ho);
    const long long bc = i / (static_cast<long long>(Wo) * Ho);
    const float* pl = in + bc * H * W;
    int best = (ho * kh) * W + wo * kw;
    float bv = pl[best];
    for (int a = 0; a < kh; ++a)
      for (int b = 0; b < kw; ++b) {
        const int j = (ho * kh + a) * W + wo * kw + b;
        const float v = pl[j];
        if (v > bv || v != v) { bv = v; best = j; }
      }
    out[i] = bv;
    idx[i] = best;
  }
}

// one thread per INPUT element: it receives its window's gradient if it is that window's argmax (windows do not overlap)
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const int* __restrict__ idx, float* __restrict__ dx, long long total, int H, int W,
                                   int Ho, int Wo, int kh, int kw) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int w = static_cast<int>(i % W), h = static_cast<int>((i / W) % H);
    const long long bc = i / (static_cast<long long>(W) * H);
    const int ho = h / kh, wo = w / kw;
    float g = 0.f;
    if (ho < Ho && wo < Wo) {
      const long long o = (bc * Ho + ho) * Wo + wo;
      if (idx[o] == h * W + w) g = dy[o];
    }
    dx[i] = g;
  }
}

// y[b][o] = sum_k x[b][k] * W[o][k] + bias[o] for a batch far smaller than the feature count (the cnn-* heads: 26 624 features
// into 12 labels for cnn-trad-pool2).  grid (Out, ceil(B / 8)): a block reads W's row once against 8 rows of x, k strided
// over its 256 threads, and reduces in a fixed order.
constexpr int LF_BT = 8;
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ W, const float* __restrict__ bias,
                                                         float* __restrict__ y, int B, int In, int Out) {
  __shared__ double sh[8];
  const int o = blockIdx.x, b0 = blockIdx.y * LF_BT;
  const float* wr = W + static_cast<long long>(o) * In;
  float acc[LF_BT];
#pragma unroll
  for (int j = 0; j < LF_BT; ++j) acc[j] = 0.f;
  for (int k = threadIdx.x; k < In; k += 256) {
    const float wv = __ldg(wr + k);
#pragma unroll
    for (int j = 0; j < LF_BT; ++j)
      if (b0 + j < B) acc[j] = fmaf(__ldg(x + static_cast<long long>(b0 + j) * In + k), wv, acc[j]);
  }
#pragma unroll
  for (int j = 0; j < LF_BT; ++j) {
    const double t = block_sum(static_cast<double>(acc[j]), sh);
    if (threadIdx.x == 0 && b0 + j < B) y[static_cast<long long>(b0 + j) * Out + o] = static_cast<float>(t + static_cast<double>(bias[o]));
  }
}

// dx[b][i] = sum_o dy[b][o] * W[o][i]
__global__ void linear_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ W, float* __restrict__ dx, int B, int In, int Out) {
  const long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (t >= static_cast<long long>(B) * In) return;
  const int i = static_cast<int>(t % In), b = static_cast<int>(t / In);
  float s = 0.f;
  for (int o = 0; o < Out; ++o) s = fmaf(dy[b * Out + o], W[static_cast<long long>(o) * In + i], s);
  dx[t] = s;
}

// dW[o][i] = sum_b dy[b][o] * x[b][i]   (i == In: the bias gradient db[o] = sum_b dy[b][o])
__global__ void linear_wgrad_full_kernel(const float* __restrict__ dy, const float* __restrict__ x, float* __restrict__ dW, float* __restrict__ db,
                                         int B, int In, int Out) {
  const long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (t >= static_cast<long long>(Out) * (In + 1)) return;
  const int i = static_cast<int>(t % (In + 1)), o = static_cast<int>(t / (In + 1));
  double s = 0.0;
  for (int b = 0; b < B; ++b) s += static_cast<double>(dy[b * Out + o]) * (i < In ? x[static_cast<long long>(b) * In + i] : 1.f);
  if (i < In) dW[static_cast<long long>(o) * In + i] = static_cast<float>(s);
  else db[o] = static_cast<float>(s);
}

// db[c] = sum over (b, pos) of dy[b][c][pos] from the partials of chan_sums_kernel(dy, nullptr)
__global__ void bias_grad_finalize_kernel(const double* __restrict__ partial, int nsplit, float* __restrict__ db, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0;
  for (int i = 0; i < nsplit; ++i) s += partial[(static_cast<long long>(c) * nsplit + i) * 2];
  db[c] = static_cast<float>(s);
}

// in-place ReLU backward for a Linear layer's activation: g = (out > 0) ? g : 0
__global__ void relu_bwd_kernel(float* __restrict__ g, const float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    if (!(out[i] > 0.f)) g[i] = 0.f;
}

__global__ void relu_fwd_kernel(float* __restrict__ x, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    x[i] = fmaxf(x[i], 0.f);
}

inline int ew_blocks(long long n) { return static_cast<int>(std::min<long long>((n + 255) / 256, 148 * 16)); }

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------------------------------------------
struct TrainParam {
  std::string name;
  int64_t numel = 0;
  float* p = nullptr;     // parameter
  float* g = nullptr;     // gradient of the last step
  float* mom = nullptr;   // SGD momentum buffer
};

struct honk_trainer {
  honk_model_config cfg{};
  int device = 0;
  int H0 = 101, W0 = 40, Hr = 101, Wr = 40;
  std::vector<TrainParam> params;            // conv0.weight, conv1.weight .. conv_n.weight, output.weight, output.bias
  std::vector<float*> running_mean, running_var, bn_mean, bn_invstd;   // [n_layers] of [C]
  float* bn_sums = nullptr;                  // [2C]
  double* partial = nullptr;                 // [C][kStatSplit][2]
  float* wflip = nullptr;                    // [C*C*9]
  std::vector<void*> owned;
  bool momentum_ready = false;               // false: the next step initialises the momentum buffers with the gradient
  int64_t steps = 0;                         // num_batches_tracked of every BatchNorm layer
  int64_t base_tracked = 0;
  // cnn-* family: geometry derived from cfg, Dropout settings
  int C1 = 0, H1 = 0, W1 = 0, H1p = 0, W1p = 0;     // conv1 output, pooled
  int C2 = 0, H2 = 0, W2 = 0, H2p = 0, W2p = 0;     // conv2 output, pooled (has_conv2)
  int flat = 0;                                    // features entering the first Linear layer
  float dropout_p = 0.5f;                          // nn.Dropout(config["dropout_prob"]), model.py:183
  uint64_t seed = 0x5eed5eedull;
  const unsigned char* user_mask[4] = {nullptr, nullptr, nullptr, nullptr};   // caller's keep masks for the NEXT step (device, uint8)
  int pi(const char* name) const {                 // index of a parameter by its state_dict key, -1 when absent
    for (size_t i = 0; i < params.size(); ++i)
      if (params[i].name == name) return static_cast<int>(i);
    return -1;
  }
};

namespace {

int t_alloc(honk_trainer* t, int64_t n_bytes, void** out) {
  void* p = nullptr;
  HONK_CUDA(cudaMalloc(&p, std::max<int64_t>(n_bytes, 4)));
  t->owned.push_back(p);
  HONK_CUDA(cudaMemset(p, 0, std::max<int64_t>(n_bytes, 4)));
  *out = p;
  return HONK_OK;
}

int t_dilation(const honk_model_config& c, int i) { return c.use_dilation ? (1 << ((i - 1) / 3)) : 1; }

struct TrainWs {
  float *y0 = nullptr, *stem = nullptr;      // relu(conv0) before the pool (pooled models) / conv1's input
  std::vector<float*> y, x;                  // per layer: relu(conv_i), BatchNorm output
  float* res[2] = {nullptr, nullptr};        // residual stream, ping-pong
  float *pooled = nullptr, *logits = nullptr, *dlogits = nullptr, *rowloss = nullptr;
  float *dx = nullptr, *dc = nullptr, *dold = nullptr, *wpart = nullptr;
  int wsplit_main = 1, wsplit_stem = 1;
};

int wgrad_bm(int Cout) { return Cout <= 32 ? 32 : 64; }

int wgrad_split(int n_cols, int Cout, long long K) {          // n_cols = Cin * KH * KW
  const int tiles = n_cols >= 128 && Cout > 32 ? ceil_div(n_cols, 128) * ceil_div(Cout, Cout > 48 ? 64 : 48)
                                                : ceil_div(n_cols, WG_BN) * ceil_div(Cout, wgrad_bm(Cout));
  long long s = ceil_div(148 * 6, tiles);
  s = std::max<long long>(1, std::min<long long>(s, std::min<long long>((K + 255) / 256, 1024)));
  return static_cast<int>(s);
}

void carve(const honk_trainer* t, int64_t B, Arena& a, TrainWs& w) {
  const honk_model_config& c = t->cfg;
  const int C = c.n_feature_maps, n = c.n_layers;
  const int64_t big = B * C * t->H0 * t->W0, map = B * C * t->Hr * t->Wr;
  const bool pooled = c.res_pool[0] > 0;
  if (pooled) w.y0 = a.take<float>(big);
  w.stem = a.take<float>(map);
  if (!pooled) w.y0 = w.stem;
  w.y.resize(n);
  w.x.resize(n);
  for (int i = 0; i < n; ++i) {
    w.y[i] = a.take<float>(map);
    w.x[i] = a.take<float>(map);
  }
  w.res[0] = a.take<float>(map);
  w.res[1] = a.take<float>(map);
  w.pooled = a.take<float>(B * C);
  w.logits = a.take<float>(B * c.n_labels);
  w.dlogits = a.take<float>(B * c.n_labels);
  w.rowloss = a.take<float>(B);
  w.dx = a.take<float>(map);
  w.dc = a.take<float>(big);
  w.dold = a.take<float>(map);
  w.wsplit_main = wgrad_split(C * 9, C, B * t->Hr * t->Wr);
  w.wsplit_stem = wgrad_split(9, C, B * t->H0 * t->W0);
  w.wpart = a.take<float>(std::max<int64_t>(static_cast<int64_t>(w.wsplit_main) * C * C * 9, static_cast<int64_t>(w.wsplit_stem) * C * 9));
}

bool wgrad_big(const ConvGeom& g) {
  static const bool off = getenv("HONK_WGRAD_SMALL_TILES") && atoi(getenv("HONK_WGRAD_SMALL_TILES")) != 0;
  return !off && g.Cin * g.KH * g.KW >= 128 && g.Cout > 32 && g.KH * g.dh < 16384 && g.KW * g.dw < 16384 && g.ph < 16384 && g.pw < 16384;
}
int wgrad_big_tm(int Cout) { return Cout > 48 ? 4 : (Cout > 32 ? 3 : 2); }

int wgrad(const float* x, const float* dy, float* part, float* dw, const ConvGeom& g, int nsplit, cudaStream_t s) {
  const long long K = static_cast<long long>(g.B) * g.Ho * g.Wo;
  const long long per = align_up((K + nsplit - 1) / nsplit, WG_BK);
  const int n = g.Cout * g.Cin * g.KH * g.KW;
  if (wgrad_big(g)) {
    const int tm = wgrad_big_tm(g.Cout);
    const dim3 grid(ceil_div(g.Cin * g.KH * g.KW, WGB_BN), nsplit, ceil_div(g.Cout, 16 * tm));
    if (tm == 4) wgrad_gemm_big_kernel<4><<<grid, WG_THREADS, 0, s>>>(x, dy, part, g, per);
    else if (tm == 3) wgrad_gemm_big_kernel<3><<<grid, WG_THREADS, 0, s>>>(x, dy, part, g, per);
    else wgrad_gemm_big_kernel<2><<<grid, WG_THREADS, 0, s>>>(x, dy, part, g, per);
  } else {
    const int bm = wgrad_bm(g.Cout);
    const dim3 grid(ceil_div(g.Cin * g.KH * g.KW, WG_BN), nsplit, ceil_div(g.Cout, bm));
    if (bm == 32) wgrad_gemm_kernel<32><<<grid, WG_THREADS, 0, s>>>(x, dy, part, g, per);
    else wgrad_gemm_kernel<64><<<grid, WG_THREADS, 0, s>>>(x, dy, part, g, per);
  }
  HONK_LAUNCH_CHECK();
  wgrad_reduce_kernel<<<ceil_div(n, 256), 256, 0, s>>>(part, dw, nsplit, n);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// cnn-* family: workspace and the step (SpeechModel.forward in training mode, utils/model.py:186-205)
// ---------------------------------------------------------------------------------------------------------------------
struct CnnWs {
  float *a1 = nullptr, *p1 = nullptr, *a2 = nullptr, *p2 = nullptr;      // dropped relu(conv) maps and their pooled versions
  int *i1 = nullptr, *i2 = nullptr;                                      // argmax of the pool windows
  unsigned char* mask[4] = {nullptr, nullptr, nullptr, nullptr};
  float *lin = nullptr, *z1 = nullptr, *e1 = nullptr, *z2 = nullptr, *e2 = nullptr;   // Linear outputs / dropped versions
  float *logits = nullptr, *dlogits = nullptr, *rowloss = nullptr;
  float *ga = nullptr, *gb = nullptr;                                    // gradient ping-pong (largest activation)
  float* wpart = nullptr;
  int split1 = 1, split2 = 1;
};

void carve_cnn(const honk_trainer* t, int64_t B, Arena& a, CnnWs& w) {
  const honk_model_config& c = t->cfg;
  const int64_t n1 = B * t->C1 * t->H1 * t->W1, n1p = B * t->C1 * t->H1p * t->W1p;
  const int64_t n2 = c.has_conv2 ? B * t->C2 * t->H2 * t->W2 : 0, n2p = c.has_conv2 ? B * t->C2 * t->H2p * t->W2p : 0;
  w.a1 = a.take<float>(n1);
  w.p1 = a.take<float>(n1p);
  w.i1 = a.take<int>(n1p);
  w.mask[0] = a.take<unsigned char>(n1);
  if (c.has_conv2) {
    w.a2 = a.take<float>(n2);
    w.p2 = a.take<float>(n2p);
    w.i2 = a.take<int>(n2p);
    w.mask[1] = a.take<unsigned char>(n2);
  }
  if (!c.tf_variant) w.lin = a.take<float>(B * 32);
  if (c.dnn1_size > 0) {
    w.z1 = a.take<float>(B * c.dnn1_size);
    w.e1 = a.take<float>(B * c.dnn1_size);
    w.mask[2] = a.take<unsigned char>(B * c.dnn1_size);
    if (c.dnn2_size > 0) {
      w.z2 = a.take<float>(B * c.dnn2_size);
      w.e2 = a.take<float>(B * c.dnn2_size);
      w.mask[3] = a.take<unsigned char>(B * c.dnn2_size);
    }
  }
  w.logits = a.take<float>(B * c.n_labels);
  w.dlogits = a.take<float>(B * c.n_labels);
  w.rowloss = a.take<float>(B);
  const int64_t big = std::max<int64_t>(std::max(n1, n2), std::max<int64_t>(B * t->flat, 1024));
  w.ga = a.take<float>(big);
  w.gb = a.take<float>(big);
  const int k1 = c.conv1_size[0] * c.conv1_size[1];
  w.split1 = wgrad_split(k1, t->C1, B * t->H1 * t->W1);
  int64_t part = static_cast<int64_t>(w.split1) * t->C1 * k1;
  if (c.has_conv2) {
    const int k2 = t->C1 * c.conv2_size[0] * c.conv2_size[1];
    w.split2 = wgrad_split(k2, t->C2, B * t->H2 * t->W2);
    part = std::max(part, static_cast<int64_t>(w.split2) * t->C2 * k2);
  }
  w.wpart = a.take<float>(part);
}

// y = x W^T + b for a Linear layer stored as parameters (wi, wi + 1)
int linear_fwd(const honk_trainer* t, int wi, const float* x, float* y, int B, cudaStream_t s) {
  const TrainParam& W = t->params[wi];
  const int Out = static_cast<int>(t->params[wi + 1].numel), In = static_cast<int>(W.numel / Out);
  linear_fwd_kernel<<<dim3(Out, ceil_div(B, LF_BT)), 256, 0, s>>>(x, W.p, t->params[wi + 1].p, y, B, In, Out);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

// gradients of a Linear layer: dW, db always; dx when asked for
int linear_bwd(honk_trainer* t, int wi, const float* x, const float* dy, float* dx, int B, cudaStream_t s) {
  TrainParam& W = t->params[wi];
  const int Out = static_cast<int>(t->params[wi + 1].numel), In = static_cast<int>(W.numel / Out);
  const long long nw = static_cast<long long>(Out) * (In + 1);
  linear_wgrad_full_kernel<<<static_cast<unsigned>((nw + 127) / 128), 128, 0, s>>>(dy, x, W.g, t->params[wi + 1].g, B, In, Out);
  HONK_LAUNCH_CHECK();
  if (dx) {
    const long long nx = static_cast<long long>(B) * In;
    linear_dgrad_kernel<<<static_cast<unsigned>((nx + 127) / 128), 128, 0, s>>>(dy, W.p, dx, B, In, Out);
    HONK_LAUNCH_CHECK();
  }
  return HONK_OK;
}

int dropout_fwd(honk_trainer* t, int site, const float* x, unsigned char* mask, float* y, long long n, cudaStream_t s) {
  const float p = t->dropout_p;
  const uint32_t seed = static_cast<uint32_t>(t->seed) ^ (static_cast<uint32_t>(t->seed >> 32) * 0x9e3779b9u) ^
                        (static_cast<uint32_t>(t->steps) * 0x85ebca6bu) ^ (static_cast<uint32_t>(site + 1) * 0xc2b2ae35u);
  dropout_fwd_kernel<<<ew_blocks(n), 256, 0, s>>>(x, t->user_mask[site], mask, y, n, p, 1.f / (1.f - p), seed);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

int cnn_step(honk_trainer* t, const float* x, const int64_t* labels, int B, float* d_scores, float* d_loss, CnnWs& w, cudaStream_t s) {
  const honk_model_config& c = t->cfg;
  const bool drop = t->dropout_p > 0.f;
  const float scale = drop ? 1.f / (1.f - t->dropout_p) : 1.f;
  const int L = c.n_labels;
  const int i_c1 = t->pi("conv1.weight"), i_c2 = t->pi("conv2.weight"), i_lin = t->pi("lin.weight"), i_d1 = t->pi("dnn1.weight"),
            i_d2 = t->pi("dnn2.weight"), i_out = t->pi("output.weight");
  const long long n1 = static_cast<long long>(B) * t->C1 * t->H1 * t->W1, n1p = static_cast<long long>(B) * t->C1 * t->H1p * t->W1p;
  const long long n2 = c.has_conv2 ? static_cast<long long>(B) * t->C2 * t->H2 * t->W2 : 0;
  const long long n2p = c.has_conv2 ? static_cast<long long>(B) * t->C2 * t->H2p * t->W2p : 0;
  ConvGeom g1{B, 1, t->H0, t->W0, t->C1, c.conv1_size[0], c.conv1_size[1], c.conv1_stride[0], c.conv1_stride[1], 0, 0, 1, 1, 0, 0};
  conv_out_dims(g1);
  ConvGeom g2{B, t->C1, t->H1p, t->W1p, t->C2, c.conv2_size[0], c.conv2_size[1], 1, 1, 0, 0, 1, 1, 0, 0};
  if (c.has_conv2) conv_out_dims(g2);
  const float* flat = nullptr;       // [B, flat] input of the first Linear layer
  const float* last = nullptr;       // input of the output layer
  // ---- forward ----------------------------------------------------------------------------------------------------
  {
    ProfSpan sp("train_fwd", s);
    HONK_TRY(conv2d_f32(x, t->params[i_c1].p, t->params[i_c1 + 1].p, w.a1, g1, true, s));
    if (drop) HONK_TRY(dropout_fwd(t, 0, w.a1, w.mask[0], w.a1, n1, s));
    maxpool_fwd_kernel<<<ew_blocks(n1p), 256, 0, s>>>(w.a1, w.p1, w.i1, n1p, t->H1, t->W1, t->H1p, t->W1p, c.conv1_pool[0], c.conv1_pool[1]);
    HONK_LAUNCH_CHECK();
    flat = w.p1;
    if (c.has_conv2) {
      HONK_TRY(conv2d_f32(w.p1, t->params[i_c2].p, t->params[i_c2 + 1].p, w.a2, g2, true, s));
      if (drop) HONK_TRY(dropout_fwd(t, 1, w.a2, w.mask[1], w.a2, n2, s));
      maxpool_fwd_kernel<<<ew_blocks(n2p), 256, 0, s>>>(w.a2, w.p2, w.i2, n2p, t->H2, t->W2, t->H2p, t->W2p, c.conv2_pool[0], c.conv2_pool[1]);
      HONK_LAUNCH_CHECK();
      flat = w.p2;
    }
    last = flat;
    if (i_lin >= 0) {
      HONK_TRY(linear_fwd(t, i_lin, last, w.lin, B, s));
      last = w.lin;
    }
    if (i_d1 >= 0) {
      HONK_TRY(linear_fwd(t, i_d1, last, w.z1, B, s));
      const long long nz = static_cast<long long>(B) * c.dnn1_size;
      if (!c.tf_variant) {
        relu_fwd_kernel<<<ew_blocks(nz), 256, 0, s>>>(w.z1, nz);
        HONK_LAUNCH_CHECK();
      }
      if (drop) HONK_TRY(dropout_fwd(t, 2, w.z1, w.mask[2], w.e1, nz, s));
      last = drop ? w.e1 : w.z1;
      if (i_d2 >= 0) {
        HONK_TRY(linear_fwd(t, i_d2, last, w.z2, B, s));
        const long long nz2 = static_cast<long long>(B) * c.dnn2_size;
        if (drop) HONK_TRY(dropout_fwd(t, 3, w.z2, w.mask[3], w.e2, nz2, s));
        last = drop ? w.e2 : w.z2;
      }
    }
    HONK_TRY(linear_fwd(t, i_out, last, w.logits, B, s));
    ce_fwd_bwd_kernel<<<ceil_div(B, 128), 128, 0, s>>>(w.logits, reinterpret_cast<const long long*>(labels), w.dlogits, w.rowloss, B, L);
    HONK_LAUNCH_CHECK();
    if (d_loss) {
      loss_mean_kernel<<<1, 256, 0, s>>>(w.rowloss, d_loss, B);
      HONK_LAUNCH_CHECK();
    }
    if (d_scores) HONK_CUDA(cudaMemcpyAsync(d_scores, w.logits, static_cast<size_t>(B) * L * 4, cudaMemcpyDeviceToDevice, s));
  }
  // ---- backward ---------------------------------------------------------------------------------------------------
  {
    ProfSpan sp("train_bwd", s);
    float* g = w.ga;                   // gradient at the input of the layer being processed
    float* g2b = w.gb;
    HONK_TRY(linear_bwd(t, i_out, last, w.dlogits, g, B, s));
    if (i_d1 >= 0) {
      if (i_d2 >= 0) {
        const long long nz2 = static_cast<long long>(B) * c.dnn2_size;
        if (drop) {
          dropout_relu_bwd_kernel<<<ew_blocks(nz2), 256, 0, s>>>(g, w.mask[3], nullptr, g, nz2, scale);
          HONK_LAUNCH_CHECK();
        }
        HONK_TRY(linear_bwd(t, i_d2, drop ? w.e1 : w.z1, g, g2b, B, s));
        std::swap(g, g2b);
      }
      const long long nz = static_cast<long long>(B) * c.dnn1_size;
      if (drop || !c.tf_variant) {
        dropout_relu_bwd_kernel<<<ew_blocks(nz), 256, 0, s>>>(g, drop ? w.mask[2] : nullptr, c.tf_variant ? nullptr : w.z1, g, nz, scale);
        HONK_LAUNCH_CHECK();
      }
      HONK_TRY(linear_bwd(t, i_d1, i_lin >= 0 ? w.lin : flat, g, g2b, B, s));
      std::swap(g, g2b);
    }
    if (i_lin >= 0) {
      HONK_TRY(linear_bwd(t, i_lin, flat, g, g2b, B, s));
      std::swap(g, g2b);
    }
    // g: gradient at the flattened pooled map
    if (c.has_conv2) {
      maxpool_bwd_kernel<<<ew_blocks(n2), 256, 0, s>>>(g, w.i2, g2b, n2, t->H2, t->W2, t->H2p, t->W2p, c.conv2_pool[0], c.conv2_pool[1]);
      HONK_LAUNCH_CHECK();
      dropout_relu_bwd_kernel<<<ew_blocks(n2), 256, 0, s>>>(g2b, drop ? w.mask[1] : nullptr, w.a2, g2b, n2, scale);
      HONK_LAUNCH_CHECK();
      std::swap(g, g2b);               // g = gradient at conv2's output (before bias / ReLU)
      chan_sums_kernel<<<dim3(t->C2, kStatSplit), 256, 0, s>>>(g, nullptr, t->partial, B, t->C2, t->H2 * t->W2);
      HONK_LAUNCH_CHECK();
      bias_grad_finalize_kernel<<<ceil_div(t->C2, 128), 128, 0, s>>>(t->partial, kStatSplit, t->params[i_c2 + 1].g, t->C2);
      HONK_LAUNCH_CHECK();
      HONK_TRY(wgrad(w.p1, g, w.wpart, t->params[i_c2].g, g2, w.split2, s));
      const int khw = g2.KH * g2.KW;
      flip_weights_kernel<<<ceil_div(t->C2 * t->C1 * khw, 256), 256, 0, s>>>(t->params[i_c2].p, t->wflip, t->C2, t->C1, khw);
      HONK_LAUNCH_CHECK();
      ConvGeom gd{B, t->C2, t->H2, t->W2, t->C1, g2.KH, g2.KW, 1, 1, g2.KH - 1, g2.KW - 1, 1, 1, 0, 0};
      conv_out_dims(gd);               // = H1p x W1p
      HONK_TRY(conv2d_f32(g, t->wflip, nullptr, g2b, gd, false, s));
      std::swap(g, g2b);               // g = gradient at the pooled conv1 map
    }
    maxpool_bwd_kernel<<<ew_blocks(n1), 256, 0, s>>>(g, w.i1, g2b, n1, t->H1, t->W1, t->H1p, t->W1p, c.conv1_pool[0], c.conv1_pool[1]);
    HONK_LAUNCH_CHECK();
    dropout_relu_bwd_kernel<<<ew_blocks(n1), 256, 0, s>>>(g2b, drop ? w.mask[0] : nullptr, w.a1, g2b, n1, scale);
    HONK_LAUNCH_CHECK();
    chan_sums_kernel<<<dim3(t->C1, kStatSplit), 256, 0, s>>>(g2b, nullptr, t->partial, B, t->C1, t->H1 * t->W1);
    HONK_LAUNCH_CHECK();
    bias_grad_finalize_kernel<<<ceil_div(t->C1, 128), 128, 0, s>>>(t->partial, kStatSplit, t->params[i_c1 + 1].g, t->C1);
    HONK_LAUNCH_CHECK();
    HONK_TRY(wgrad(x, g2b, w.wpart, t->params[i_c1].g, g1, w.split1, s));
  }
  return HONK_OK;
}

// torch.optim.SGD.step over every parameter of the handle
int sgd_apply(honk_trainer* t, const honk_sgd* opt, cudaStream_t s) {
  ProfSpan sp("train_sgd", s);
  const int first = t->momentum_ready ? 0 : 1;
  for (TrainParam& p : t->params) {
    sgd_kernel<<<ceil_div(static_cast<int>(p.numel), 256), 256, 0, s>>>(p.p, p.g, p.mom, static_cast<int>(p.numel), opt->lr, opt->momentum,
                                                                       opt->weight_decay, opt->nesterov, first);
    HONK_LAUNCH_CHECK();
  }
  if (opt->momentum != 0.f) t->momentum_ready = true;
  t->steps += 1;
  return HONK_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------------
extern "C" int honk_trainer_create(const honk_model_config* cfg, const honk_tensor* tensors, int32_t n_tensors, int device,
                                   honk_trainer** out) {
  HONK_CHECK(cfg && out && tensors, HONK_ERR_INVALID, "honk_trainer_create: null argument");
  HONK_CHECK(cfg->family == HONK_FAMILY_RES || cfg->family == HONK_FAMILY_CNN, HONK_ERR_INVALID, "honk_trainer_create: unknown family");
  // the strict state_dict check (names, shapes) is the model handle's: build one and take its validated host copy
  honk_model* m = nullptr;
  HONK_TRY(honk_model_create(cfg, tensors, n_tensors, device, &m));
  std::unique_ptr<honk_model, int (*)(honk_model*)> guard(m, honk_model_destroy);
  HONK_CUDA(cudaSetDevice(device));
  std::unique_ptr<honk_trainer> t(new honk_trainer());
  t->cfg = *cfg;
  t->device = device;
  t->H0 = m->H0; t->W0 = m->W0; t->Hr = m->Hr; t->Wr = m->Wr;
  auto fail = [&](int rc) {
    for (void* p : t->owned) cudaFree(p);
    return rc;
  };
  auto add_param = [&](const std::string& name) -> int {
    const HostTensor& h = m->host.at(name);
    TrainParam p;
    p.name = name;
    p.numel = h.numel();
    HONK_TRY(t_alloc(t.get(), p.numel * 4, reinterpret_cast<void**>(&p.p)));
    HONK_TRY(t_alloc(t.get(), p.numel * 4, reinterpret_cast<void**>(&p.g)));
    HONK_TRY(t_alloc(t.get(), p.numel * 4, reinterpret_cast<void**>(&p.mom)));
    HONK_CUDA(cudaMemcpy(p.p, h.data.data(), p.numel * 4, cudaMemcpyHostToDevice));
    t->params.push_back(p);
    return HONK_OK;
  };
  int rc = HONK_OK;
  if (cfg->family == HONK_FAMILY_CNN) {
    const honk_model_config& c = *cfg;
    HONK_CHECK(!c.has_conv2 || (c.conv2_stride[0] == 1 && c.conv2_stride[1] == 1), HONK_ERR_UNSUPPORTED,
               "honk_trainer_create: conv2 with a stride is not supported (no reference config has one)");
    t->H0 = c.height; t->W0 = c.width;
    t->C1 = c.n_feature_maps1;
    t->H1 = (c.height - c.conv1_size[0]) / c.conv1_stride[0] + 1;
    t->W1 = (c.width - c.conv1_size[1]) / c.conv1_stride[1] + 1;
    t->H1p = t->H1 / c.conv1_pool[0]; t->W1p = t->W1 / c.conv1_pool[1];
    t->flat = t->C1 * t->H1p * t->W1p;
    if (c.has_conv2) {
      t->C2 = c.n_feature_maps2;
      t->H2 = t->H1p - c.conv2_size[0] + 1; t->W2 = t->W1p - c.conv2_size[1] + 1;
      t->H2p = t->H2 / c.conv2_pool[0]; t->W2p = t->W2 / c.conv2_pool[1];
      t->flat = t->C2 * t->H2p * t->W2p;
    }
    const char* names[] = {"conv1", "conv2", "lin", "dnn1", "dnn2", "output"};
    for (const char* nm : names) {
      if (!m->host.count(std::string(nm) + ".weight")) continue;
      if (rc == HONK_OK) rc = add_param(std::string(nm) + ".weight");
      if (rc == HONK_OK) rc = add_param(std::string(nm) + ".bias");
    }
    if (rc != HONK_OK) return fail(rc);
    int64_t max_conv_w = t->params[0].numel;
    if (c.has_conv2) max_conv_w = std::max(max_conv_w, t->params[t->pi("conv2.weight")].numel);
    const int maxC = std::max(t->C1, t->C2);
    if (rc == HONK_OK) rc = t_alloc(t.get(), static_cast<int64_t>(maxC) * kStatSplit * 2 * 8, reinterpret_cast<void**>(&t->partial));
    if (rc == HONK_OK) rc = t_alloc(t.get(), max_conv_w * 4, reinterpret_cast<void**>(&t->wflip));
    if (rc != HONK_OK) return fail(rc);
    *out = t.release();
    return HONK_OK;
  }
  const int C = cfg->n_feature_maps, n = cfg->n_layers;
  rc = add_param("conv0.weight");
  for (int i = 1; rc == HONK_OK && i <= n; ++i) rc = add_param("conv" + std::to_string(i) + ".weight");
  if (rc == HONK_OK) rc = add_param("output.weight");
  if (rc == HONK_OK) rc = add_param("output.bias");
  if (rc != HONK_OK) return fail(rc);
  auto upload_stat = [&](const std::string& name, float** d) -> int {
    const HostTensor& h = m->host.at(name);
    HONK_TRY(t_alloc(t.get(), C * 4, reinterpret_cast<void**>(d)));
    HONK_CUDA(cudaMemcpy(*d, h.data.data(), C * 4, cudaMemcpyHostToDevice));
    return HONK_OK;
  };
  t->running_mean.resize(n); t->running_var.resize(n); t->bn_mean.resize(n); t->bn_invstd.resize(n);
  for (int i = 0; rc == HONK_OK && i < n; ++i) {
    const std::string s = "bn" + std::to_string(i + 1);
    rc = upload_stat(s + ".running_mean", &t->running_mean[i]);
    if (rc == HONK_OK) rc = upload_stat(s + ".running_var", &t->running_var[i]);
    if (rc == HONK_OK) rc = t_alloc(t.get(), C * 4, reinterpret_cast<void**>(&t->bn_mean[i]));
    if (rc == HONK_OK) rc = t_alloc(t.get(), C * 4, reinterpret_cast<void**>(&t->bn_invstd[i]));
  }
  if (rc == HONK_OK) rc = t_alloc(t.get(), 2 * C * 4, reinterpret_cast<void**>(&t->bn_sums));
  if (rc == HONK_OK) rc = t_alloc(t.get(), static_cast<int64_t>(C) * kStatSplit * 2 * 8, reinterpret_cast<void**>(&t->partial));
  if (rc == HONK_OK) rc = t_alloc(t.get(), static_cast<int64_t>(C) * C * 9 * 4, reinterpret_cast<void**>(&t->wflip));
  if (rc != HONK_OK) return fail(rc);
  *out = t.release();
  return HONK_OK;
}

extern "C" int honk_trainer_destroy(honk_trainer* t) {
  if (!t) return HONK_OK;
  cudaSetDevice(t->device);
  for (void* p : t->owned) cudaFree(p);
  delete t;
  return HONK_OK;
}

extern "C" int64_t honk_trainer_workspace_bytes(const honk_trainer* t, int64_t n_clips) {
  if (!t || n_clips <= 0) return 0;
  Arena a(nullptr, 0);
  if (t->cfg.family == HONK_FAMILY_CNN) {
    CnnWs w;
    carve_cnn(t, n_clips, a, w);
  } else {
    TrainWs w;
    carve(t, n_clips, a, w);
  }
  return align_up(a.off, 256);
}

extern "C" int honk_trainer_reset_optimizer(honk_trainer* t) {
  HONK_CHECK(t, HONK_ERR_INVALID, "honk_trainer_reset_optimizer: null handle");
  t->momentum_ready = false;
  return HONK_OK;
}

extern "C" int64_t honk_trainer_steps(const honk_trainer* t) { return t ? t->steps : 0; }

extern "C" int honk_trainer_set_dropout(honk_trainer* t, float prob, uint64_t seed) {
  HONK_CHECK(t, HONK_ERR_INVALID, "honk_trainer_set_dropout: null handle");
  HONK_CHECK(prob >= 0.f && prob < 1.f, HONK_ERR_INVALID, "honk_trainer_set_dropout: probability %g outside [0, 1)", prob);
  t->dropout_p = prob;
  t->seed = seed;
  return HONK_OK;
}

extern "C" int honk_trainer_set_dropout_masks(honk_trainer* t, const unsigned char* const* d_masks, int32_t n_masks) {
  HONK_CHECK(t && (d_masks || n_masks == 0) && n_masks >= 0 && n_masks <= 4, HONK_ERR_INVALID, "honk_trainer_set_dropout_masks: bad argument");
  for (int i = 0; i < 4; ++i) t->user_mask[i] = i < n_masks ? d_masks[i] : nullptr;
  return HONK_OK;
}

// name: a state_dict key ("conv3.weight", "bn2.running_var", "output.bias"), or "grad:<parameter>" / "momentum:<parameter>".
extern "C" int honk_trainer_read(const honk_trainer* t, const char* name, float* h_out, int64_t numel) {
  HONK_CHECK(t && name && h_out, HONK_ERR_INVALID, "honk_trainer_read: null argument");
  HONK_CUDA(cudaSetDevice(t->device));
  std::string key(name);
  int what = 0;
  if (key.rfind("grad:", 0) == 0) { what = 1; key = key.substr(5); }
  else if (key.rfind("momentum:", 0) == 0) { what = 2; key = key.substr(9); }
  const float* src = nullptr;
  int64_t n = 0;
  for (const TrainParam& p : t->params)
    if (p.name == key) { src = what == 0 ? p.p : what == 1 ? p.g : p.mom; n = p.numel; }
  if (!src && what == 0) {
    for (int i = 0; i < t->cfg.n_layers; ++i) {
      const std::string s = "bn" + std::to_string(i + 1);
      if (key == s + ".running_mean") { src = t->running_mean[i]; n = t->cfg.n_feature_maps; }
      if (key == s + ".running_var") { src = t->running_var[i]; n = t->cfg.n_feature_maps; }
    }
  }
  HONK_CHECK(src, HONK_ERR_INVALID, "honk_trainer_read: unknown tensor '%s'", name);
  HONK_CHECK(n == numel, HONK_ERR_INVALID, "honk_trainer_read: '%s' has %lld elements, caller expects %lld", name,
             static_cast<long long>(n), static_cast<long long>(numel));
  HONK_CUDA(cudaDeviceSynchronize());
  HONK_CUDA(cudaMemcpy(h_out, src, n * 4, cudaMemcpyDeviceToHost));
  return HONK_OK;
}

// One optimisation step (train.py:127-139).  d_scores [n_clips, n_labels] receives the TRAINING-mode scores the reference
// prints its "train step" accuracy from; d_loss one float, the mean CrossEntropyLoss.  Either may be nullptr.
extern "C" int honk_trainer_step(honk_trainer* t, const float* d_mfcc, const int64_t* d_labels, int64_t n_clips, int32_t height,
                                 int32_t width, const honk_sgd* opt, float* d_scores, float* d_loss, void* d_workspace,
                                 int64_t workspace_bytes, void* stream) {
  HONK_CHECK(t && d_mfcc && d_labels && opt, HONK_ERR_INVALID, "honk_trainer_step: null argument");
  HONK_CHECK(height == t->H0 && width == t->W0, HONK_ERR_INVALID, "honk_trainer_step: clips are %d x %d, the model takes %d x %d",
             height, width, t->H0, t->W0);
  HONK_CHECK(n_clips >= (t->cfg.family == HONK_FAMILY_RES ? 2 : 1) && n_clips <= 65536, HONK_ERR_INVALID,
             "honk_trainer_step: batch of %lld clips (BatchNorm in training mode needs at least 2; at most 65536)",
             static_cast<long long>(n_clips));
  HONK_CHECK(opt->momentum >= 0.f && opt->weight_decay >= 0.f && (!opt->nesterov || opt->momentum > 0.f), HONK_ERR_INVALID,
             "honk_trainer_step: bad SGD settings (Nesterov needs a momentum)");
  HONK_CUDA(cudaSetDevice(t->device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const honk_model_config& c = t->cfg;
  if (c.family == HONK_FAMILY_CNN) {
    Arena a(d_workspace, workspace_bytes);
    CnnWs w;
    carve_cnn(t, n_clips, a, w);
    HONK_CHECK(d_workspace && a.ok(), HONK_ERR_WORKSPACE, "honk_trainer_step: workspace of %lld bytes, %lld needed",
               static_cast<long long>(workspace_bytes), static_cast<long long>(align_up(a.off, 256)));
    const int rc = cnn_step(t, d_mfcc, d_labels, static_cast<int>(n_clips), d_scores, d_loss, w, s);
    for (auto& um : t->user_mask) um = nullptr;                   // the caller's masks serve one step
    HONK_TRY(rc);
    return sgd_apply(t, opt, s);
  }
  const int B = static_cast<int>(n_clips), C = c.n_feature_maps, n = c.n_layers, L = c.n_labels;
  const int H0 = t->H0, W0 = t->W0, Hr = t->Hr, Wr = t->Wr, HW = Hr * Wr;
  const bool pooled = c.res_pool[0] > 0;
  Arena a(d_workspace, workspace_bytes);
  TrainWs w;
  carve(t, n_clips, a, w);
  HONK_CHECK(d_workspace && a.ok(), HONK_ERR_WORKSPACE, "honk_trainer_step: workspace of %lld bytes, %lld needed",
             static_cast<long long>(workspace_bytes), static_cast<long long>(align_up(a.off, 256)));
  const long long map = static_cast<long long>(B) * C * HW, big = static_cast<long long>(B) * C * H0 * W0;
  const long long count = static_cast<long long>(B) * HW;
  TrainParam& P0 = t->params[0];
  TrainParam& Pw = t->params[n + 1];
  TrainParam& Pb = t->params[n + 2];

  // ---- forward, training mode -----------------------------------------------------------------------------------
  {
    ProfSpan sp("train_fwd", s);
    ConvGeom g0{B, 1, H0, W0, C, 3, 3, 1, 1, 1, 1, 1, 1, 0, 0};
    conv_out_dims(g0);
    HONK_TRY(conv2d_f32(d_mfcc, P0.p, nullptr, w.y0, g0, true, s));
    if (pooled) HONK_TRY(pool2d_f32(w.y0, w.stem, B * C, H0, W0, c.res_pool[0], c.res_pool[1], false, s));
    const float* cur = w.stem;
    const float* old = w.stem;
    int flip = 0;
    for (int i = 1; i <= n; ++i) {
      const int d = t_dilation(c, i);
      ConvGeom gi{B, C, Hr, Wr, C, 3, 3, 1, 1, d, d, d, d, 0, 0};
      conv_out_dims(gi);
      HONK_TRY(conv2d_f32(cur, t->params[i].p, nullptr, w.y[i - 1], gi, true, s));
      const float* r = w.y[i - 1];
      if (i % 2 == 0) {
        add_f32_kernel<<<ew_blocks(map), 256, 0, s>>>(w.y[i - 1], old, w.res[flip], map);
        HONK_LAUNCH_CHECK();
        r = old = w.res[flip];
        flip ^= 1;
      }
      chan_sums_kernel<<<dim3(C, kStatSplit), 256, 0, s>>>(r, nullptr, t->partial, B, C, HW);
      HONK_LAUNCH_CHECK();
      bn_train_finalize_kernel<<<ceil_div(C, 128), 128, 0, s>>>(t->partial, kStatSplit, count, t->bn_mean[i - 1], t->bn_invstd[i - 1],
                                                               t->running_mean[i - 1], t->running_var[i - 1], C);
      HONK_LAUNCH_CHECK();
      bn_apply_kernel<<<ew_blocks(map), 256, 0, s>>>(r, t->bn_mean[i - 1], t->bn_invstd[i - 1], w.x[i - 1], C, HW, map);
      HONK_LAUNCH_CHECK();
      cur = w.x[i - 1];
    }
    HONK_TRY(mean_hw_f32(cur, w.pooled, B * C, HW, s));
    ConvGeom gl{B, C, 1, 1, L, 1, 1, 1, 1, 0, 0, 1, 1, 1, 1};
    HONK_TRY(conv2d_f32(w.pooled, Pw.p, Pb.p, w.logits, gl, false, s));
    ce_fwd_bwd_kernel<<<ceil_div(B, 128), 128, 0, s>>>(w.logits, reinterpret_cast<const long long*>(d_labels), w.dlogits, w.rowloss, B, L);
    HONK_LAUNCH_CHECK();
    if (d_loss) {
      loss_mean_kernel<<<1, 256, 0, s>>>(w.rowloss, d_loss, B);
      HONK_LAUNCH_CHECK();
    }
    if (d_scores) HONK_CUDA(cudaMemcpyAsync(d_scores, w.logits, static_cast<size_t>(B) * L * 4, cudaMemcpyDeviceToDevice, s));
  }

  // ---- backward -------------------------------------------------------------------------------------------------
  {
    ProfSpan sp("train_bwd", s);
    linear_wgrad_kernel<<<ceil_div(L * (C + 1), 128), 128, 0, s>>>(w.dlogits, w.pooled, Pw.g, Pb.g, B, L, C);
    HONK_LAUNCH_CHECK();
    head_dgrad_kernel<<<B * C, 128, 0, s>>>(w.dlogits, Pw.p, w.dx, L, C, HW);
    HONK_LAUNCH_CHECK();
    bool have_dold = false;
    for (int i = n; i >= 1; --i) {
      const int d = t_dilation(c, i);
      const float* xin = i == 1 ? w.stem : w.x[i - 2];
      chan_sums_kernel<<<dim3(C, kStatSplit), 256, 0, s>>>(w.dx, w.x[i - 1], t->partial, B, C, HW);
      HONK_LAUNCH_CHECK();
      bn_bwd_finalize_kernel<<<ceil_div(C, 128), 128, 0, s>>>(t->partial, kStatSplit, count, t->bn_sums, C);
      HONK_LAUNCH_CHECK();
      const bool even = i % 2 == 0;
      bn_relu_bwd_kernel<<<ew_blocks(map), 256, 0, s>>>(w.dx, w.x[i - 1], w.y[i - 1], t->bn_sums, t->bn_invstd[i - 1],
                                                       (even && have_dold) ? w.dold : nullptr, even ? w.dold : nullptr, w.dc, C, HW, map);
      HONK_LAUNCH_CHECK();
      if (even) have_dold = true;
      ConvGeom gi{B, C, Hr, Wr, C, 3, 3, 1, 1, d, d, d, d, 0, 0};
      conv_out_dims(gi);
      HONK_TRY(wgrad(xin, w.dc, w.wpart, t->params[i].g, gi, w.wsplit_main, s));
      flip_weights_kernel<<<ceil_div(C * C * 9, 256), 256, 0, s>>>(t->params[i].p, t->wflip, C, C, 9);
      HONK_LAUNCH_CHECK();
      HONK_TRY(conv2d_f32(w.dc, t->wflip, nullptr, w.dx, gi, false, s));   // dx now holds the gradient at conv_i's input
    }
    stem_bwd_kernel<<<ew_blocks(big), 256, 0, s>>>(w.dx, have_dold ? w.dold : nullptr, w.y0, w.dc, H0, W0, pooled ? c.res_pool[0] : 1,
                                                  pooled ? c.res_pool[1] : 1, Hr, Wr, big);
    HONK_LAUNCH_CHECK();
    {
      ConvGeom g0{B, 1, H0, W0, C, 3, 3, 1, 1, 1, 1, 1, 1, 0, 0};
      conv_out_dims(g0);
      HONK_TRY(wgrad(d_mfcc, w.dc, w.wpart, P0.g, g0, w.wsplit_stem, s));
    }
  }

  HONK_TRY(sgd_apply(t, opt, s));
  return HONK_OK;
}
