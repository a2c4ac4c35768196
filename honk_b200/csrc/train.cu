// One SGD training step of the res* family on the GPU (reference: utils/train.py:127-139 - model.train(), forward,
// CrossEntropyLoss, backward, torch.optim.SGD.step - over SpeechResModel, utils/model.py:82-121).
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
                                   const float* __restrict__ dold_in, float* __restrict__ dold_out, float* __restrict__ dc,
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

// wt[ci][co][kh][kw] = w[co][ci][2-kh][2-kw]: the data gradient of a 3x3 stride-1 convolution with pad == dilation is
// the same convolution of dy with these weights.
__global__ void flip_weights_kernel(const float* __restrict__ w, float* __restrict__ wt, int Cout, int Cin) {
  const int n = Cout * Cin * 9;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int tap = i % 9, ci = (i / 9) % Cin, co = i / (9 * Cin);
    wt[(ci * Cout + co) * 9 + (8 - tap)] = w[i];
  }
}
