// tcgen05 path of the res* models: plan (weight packing, unit tables), conv0 and tail kernels,
// TMA tensor maps and the per-chunk forward.  The dilated 3x3 layers are in res_tc_conv.cu.
#include "res_tc.cuh"

#include <cudaTypedefs.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "model.cuh"

namespace honk {

struct ResTcPlan {
  int ncg = 6, N = 48, C = 45;
  int H0 = 101, W0 = 40, H = 101, W = 40;     // input size, feature-map size after the optional pool
  int pool_h = 0, pool_w = 0;
  int n_layers = 0, n_labels = 12;
  int chunk = 74;                              // clips per pass (activations stay L2-resident)
  std::vector<TcLayerGeom> geom;               // per conv layer 1..n
  std::vector<__half*> wpack;                  // device, per layer
  std::vector<float*> bn_scale, bn_shift;      // device [N], per layer
  float* w0 = nullptr;                         // conv0 weights [9][N] float
  float* wout = nullptr;                       // output.weight [L][N] float (padded)
  float* bout = nullptr;
  std::vector<void*> owned;
  PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  int grid_cap = 148;
};

namespace {

// ---- conv0: 1 -> C, 3x3, pad 1, ReLU, optional AvgPool, FP32 on CUDA cores ---------------------
// (model.py:107-111).  One CTA per clip; writes the planar f16 layout consumed by the UMMA layers.
template <int NCG>
__global__ void __launch_bounds__(256) tc_conv0_kernel(const float* __restrict__ mfcc, const float* __restrict__ w0,
                                                      __half* __restrict__ out, int H0, int W0, int ph, int pw,
                                                      int H, int W) {
  constexpr int N = NCG * 8;
  extern __shared__ float sm0[];
  float* xs = sm0;                              // [(H0+2)][(W0+2)] zero-bordered input
  float* ws = sm0 + (H0 + 2) * (W0 + 2);        // [9][N]
  const int clip = blockIdx.x, tid = threadIdx.x;
  const int Wp = W0 + 2;
  for (int i = tid; i < (H0 + 2) * Wp; i += blockDim.x) {
    const int y = i / Wp - 1, x = i % Wp - 1;
    xs[i] = (y >= 0 && y < H0 && x >= 0 && x < W0) ? __ldg(mfcc + (static_cast<long long>(clip) * H0 + y) * W0 + x) : 0.f;
  }
  for (int i = tid; i < 9 * N; i += blockDim.x) ws[i] = __ldg(w0 + i);
  __syncthreads();
  const int pool = ph > 0;
  const int kh = pool ? ph : 1, kw = pool ? pw : 1;
  const float inv = 1.f / static_cast<float>(kh * kw);
  const int items = NCG * W * H;
  for (int idx = tid; idx < items; idx += blockDim.x) {
    const int h = idx % H, w = (idx / H) % W, cg = idx / (H * W);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int py = 0; py < kh; ++py)
      for (int px = 0; px < kw; ++px) {
        const int y = h * kh + py, x = w * kw + px;          // conv output position (time, coeff)
        float t[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = 0.f;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const float v = xs[(y + tap / 3) * Wp + x + tap % 3];
#pragma unroll
          for (int k = 0; k < 8; ++k) t[k] = fmaf(v, ws[tap * N + cg * 8 + k], t[k]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += fmaxf(t[k], 0.f);
      }
    __half2 o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = __floats2half2_rn(acc[2 * k] * inv, acc[2 * k + 1] * inv);
    *reinterpret_cast<uint4*>(out + ((static_cast<long long>(clip) * NCG + cg) * W * H + static_cast<long long>(w) * H + h) * 8) =
        *reinterpret_cast<uint4*>(o);
  }
}

// ---- tail: logits = W_o * (pool / HW) + b_o ; the last BatchNorm is already inside `pool` ------
__global__ void tc_tail_kernel(const float* __restrict__ pool, const float* __restrict__ wout,
                               const float* __restrict__ bout, float* __restrict__ logits, int n_clips, int N,
                               int L, float inv_hw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_clips * L) return;
  const int clip = i / L, l = i - clip * L;
  float acc = 0.f;
  for (int k = 0; k < N; ++k) acc = fmaf(pool[clip * N + k] * inv_hw, __ldg(wout + l * N + k), acc);
  logits[i] = acc + __ldg(bout + l);
}

template <typename T>
int dev_upload(ResTcPlan* p, const std::vector<T>& h, T** d) {
  void* ptr = nullptr;
  HONK_CUDA(cudaMalloc(&ptr, std::max<size_t>(h.size(), 1) * sizeof(T)));
  p->owned.push_back(ptr);
  HONK_CUDA(cudaMemcpy(ptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  *d = static_cast<T*>(ptr);
  return HONK_OK;
}

// Splits the W/d columns of each residue chain into segments; picks the segment size that
// minimises the number of 128-position tiles subject to the TMEM (<= max_tiles) and SMEM limits.
void plan_layer(TcLayerGeom& g, int H, int W, int d, int ncg, int max_smem) {
  const int N = ncg * 8, wbytes = 9 * ncg * N * 16, nslot = 512 / N;
  g.H = H; g.W = W; g.d = d;
  g.strided = (d <= 8) ? 1 : 0;              // TMA element strides are limited to 8
  // per-column boxes land at multiples of Hp*16 bytes, which TMA wants 128-byte aligned
  g.Hp = g.strided ? H + d : static_cast<int>(align_up(H + d, 8));
  int best_seg = 1;
  long long best_cost = -1;
  const int max_chain = (W + d - 1) / d;
  for (int seg = 1; seg <= max_chain; ++seg) {
    if ((seg + 2) * d > 256 && g.strided) break;                       // TMA box limit
    const int tiles_full = (seg * g.Hp + 127) / 128;
    if (tiles_full > nslot - 2) break;
    const int stage = (seg + 2) * g.Hp * 32;
    if (2 * (stage + 512) + wbytes + 2048 > max_smem) break;
    long long tiles = 0, units = 0;
    for (int r = 0; r < std::min(d, W); ++r) {
      const int len = (W - r + d - 1) / d;
      for (int s0 = 0; s0 < len; s0 += seg) {
        tiles += (std::min(seg, len - s0) * g.Hp + 127) / 128;
        ++units;
      }
    }
    if (units > TC_MAX_UNITS) continue;
    const long long cost = tiles * 1000 + units * 60;                  // small per-unit overhead
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_seg = seg; }
  }
  g.box_cols = best_seg + 2;
  g.n_units = 0;
  for (int r = 0; r < std::min(d, W); ++r) {
    const int len = (W - r + d - 1) / d;
    for (int s0 = 0; s0 < len; s0 += best_seg) {
      TcUnit u;
      u.n_cols = static_cast<short>(std::min(best_seg, len - s0));
      u.w_first = static_cast<short>(r + (s0 - 1) * d);
      u.n_tiles = static_cast<short>((u.n_cols * g.Hp + 127) / 128);
      u.pad = 0;
      g.units[g.n_units++] = u;
    }
  }
  g.stage_bytes = g.box_cols * g.Hp * 32;
  g.stage_stride = static_cast<int>(align_up(g.stage_bytes + 512, 128));
  const int avail = max_smem - wbytes - 1024 - 1024;
  g.n_stages = std::max(2, std::min(6, avail / g.stage_stride));
  g.smem_bytes = std::max(1024 + g.n_stages * g.stage_stride + wbytes + 1024, 120 * 1024);   // > half an SM: 1 CTA/SM
}

}  // namespace

void res_tc_plan_free(ResTcPlan* p) {
  if (!p) return;
  for (void* q : p->owned) cudaFree(q);
  delete p;
}

int res_tc_plan_create(honk_model* m) {
  m->tc = nullptr;
  const honk_model_config& c = m->cfg;
  const int C = c.n_feature_maps;
  if (c.family != HONK_FAMILY_RES || C > 48) return HONK_OK;            // wider nets stay on the FP32 path
  if (const char* e = getenv("HONK_DISABLE_TC")) if (atoi(e)) return HONK_OK;
  std::unique_ptr<ResTcPlan, void (*)(ResTcPlan*)> p(new ResTcPlan(), res_tc_plan_free);
  p->C = C;
  p->ncg = C <= 32 ? 4 : 6;
  p->N = p->ncg * 8;
  p->H0 = m->H0; p->W0 = m->W0; p->H = m->Hr; p->W = m->Wr;
  p->pool_h = c.res_pool[0]; p->pool_w = c.res_pool[1];
  p->n_layers = c.n_layers;
  p->n_labels = c.n_labels;
  if (p->H + 16 > 256 || (p->H0 + 2) * (p->W0 + 2) * 4 + 9 * p->N * 4 > 200 * 1024) return HONK_OK;
  if (const char* e = getenv("HONK_TC_CHUNK")) p->chunk = std::max(1, atoi(e));
  p->grid_cap = m->n_sms;
  int max_smem = 0;
  HONK_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device));
  max_smem -= 4096;                                                      // static shared + margin
  cudaDriverEntryPointQueryResult qres;
  void* fn = nullptr;
  HONK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  HONK_CHECK(fn && qres == cudaDriverEntryPointSuccess, HONK_ERR_CUDA, "cuTensorMapEncodeTiled unavailable");
  p->encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  HONK_TRY(tc_conv_prepare(p->ncg, max_smem));
  const int N = p->N, ncg = p->ncg;

  // conv0 weights [tap][N]
  {
    const HostTensor& w = m->host.at("conv0.weight");
    std::vector<float> h(9 * N, 0.f);
    for (int co = 0; co < C; ++co)
      for (int t = 0; t < 9; ++t) h[t * N + co] = w.data[co * 9 + t];
    HONK_TRY(dev_upload(p.get(), h, &p->w0));
    const int smem0 = ((p->H0 + 2) * (p->W0 + 2) + 9 * N) * 4;
    if (ncg == 6) HONK_CUDA(cudaFuncSetAttribute(tc_conv0_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem0));
    else HONK_CUDA(cudaFuncSetAttribute(tc_conv0_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem0));
  }
  for (int i = 1; i <= c.n_layers; ++i) {
    const std::string si = std::to_string(i);
    const HostTensor& w = m->host.at("conv" + si + ".weight");          // [C][C][3][3]
    // B operand planes: [tap][cin group][cout][cin % 8], zero padded to N x N
    std::vector<__half> hw(static_cast<size_t>(9) * ncg * N * 8, __float2half(0.f));
    for (int co = 0; co < C; ++co)
      for (int ci = 0; ci < C; ++ci)
        for (int t = 0; t < 9; ++t)
          hw[((static_cast<size_t>(t) * ncg + ci / 8) * N + co) * 8 + ci % 8] =
              __float2half(w.data[(static_cast<size_t>(co) * C + ci) * 9 + t]);
    __half* dw = nullptr;
    HONK_TRY(dev_upload(p.get(), hw, &dw));
    p->wpack.push_back(dw);
    const HostTensor& mean = m->host.at("bn" + si + ".running_mean");
    const HostTensor& var = m->host.at("bn" + si + ".running_var");
    std::vector<float> sc(N, 0.f), sh(N, 0.f);
    for (int k = 0; k < C; ++k) {
      sc[k] = 1.0f / std::sqrt(var.data[k] + 1e-5f);                     // BatchNorm2d(affine=False), eval
      sh[k] = -mean.data[k] * sc[k];
    }
    float *dsc = nullptr, *dsh = nullptr;
    HONK_TRY(dev_upload(p.get(), sc, &dsc));
    HONK_TRY(dev_upload(p.get(), sh, &dsh));
    p->bn_scale.push_back(dsc);
    p->bn_shift.push_back(dsh);
    TcLayerGeom g{};
    const int d = c.use_dilation ? (1 << ((i - 1) / 3)) : 1;
    plan_layer(g, p->H, p->W, d, ncg, max_smem);
    HONK_CHECK(g.n_units >= 1 && g.smem_bytes <= max_smem, HONK_ERR_UNSUPPORTED, "tcgen05 plan does not fit (layer %d)", i);
    p->geom.push_back(g);
  }
  {
    const HostTensor& w = m->host.at("output.weight");                   // [L][C]
    std::vector<float> h(static_cast<size_t>(c.n_labels) * N, 0.f);
    for (int l = 0; l < c.n_labels; ++l)
      for (int k = 0; k < C; ++k) h[static_cast<size_t>(l) * N + k] = w.data[static_cast<size_t>(l) * C + k];
    HONK_TRY(dev_upload(p.get(), h, &p->wout));
    HONK_TRY(dev_upload(p.get(), m->host.at("output.bias").data, &p->bout));
  }
  m->tc = p.release();
  return HONK_OK;
}

int64_t res_tc_workspace_bytes(const honk_model* m, int64_t n_clips) {
  const ResTcPlan* p = m->tc;
  const int64_t nb = std::min<int64_t>(n_clips, p->chunk);
  Arena a(nullptr, 0);
  for (int i = 0; i < 3; ++i) a.take<__half>(nb * p->ncg * p->W * p->H * 8);
  a.take<float>(nb * p->N);
  return align_up(a.off, 256);
}

static int encode_map(const ResTcPlan* p, const TcLayerGeom& g, const __half* buf, int nb, CUtensorMap* out) {
  const cuuint64_t dims[5] = {8, static_cast<cuuint64_t>(g.H), static_cast<cuuint64_t>(g.W),
                              static_cast<cuuint64_t>(p->ncg), static_cast<cuuint64_t>(nb)};
  const cuuint64_t strides[4] = {16, 16ull * g.H, 16ull * g.H * g.W, 16ull * g.H * g.W * p->ncg};
  cuuint32_t box[5] = {8, static_cast<cuuint32_t>(g.Hp), 1, 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  if (g.strided) {
    box[2] = static_cast<cuuint32_t>(g.box_cols * g.d);
    box[3] = 2;
    estr[2] = static_cast<cuuint32_t>(g.d);
  }
  const CUresult r = p->encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, const_cast<__half*>(buf), dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HONK_CHECK(r == CUDA_SUCCESS, HONK_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) H=%d W=%d d=%d box=%u", static_cast<int>(r),
             g.H, g.W, g.d, box[2]);
  return HONK_OK;
}

int res_tc_forward(const honk_model* m, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws,
                   int64_t ws_bytes, cudaStream_t s) {
  const ResTcPlan* p = m->tc;
  const int nbmax = static_cast<int>(std::min<int64_t>(n_clips, p->chunk));
  Arena ar(ws, ws_bytes);
  const int64_t per = static_cast<int64_t>(p->ncg) * p->W * p->H * 8;
  __half* bufA = ar.take<__half>(nbmax * per);     // conv0 output, then the raw residual stream
  __half* bufB = ar.take<__half>(nbmax * per);
  __half* bufC = ar.take<__half>(nbmax * per);
  float* pool = ar.take<float>(static_cast<int64_t>(nbmax) * p->N);
  HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: tcgen05 workspace overrun");
  std::vector<CUtensorMap> maps(p->n_layers);
  for (int i = 1; i <= p->n_layers; ++i) {
    const __half* in = (i == 1) ? bufA : ((i % 2 == 0) ? bufB : bufC);
    HONK_TRY(encode_map(p, p->geom[i - 1], in, nbmax, &maps[i - 1]));
  }
  const int smem0 = ((p->H0 + 2) * (p->W0 + 2) + 9 * p->N) * 4;
  for (int64_t c0 = 0; c0 < n_clips; c0 += p->chunk) {
    const int nb = static_cast<int>(std::min<int64_t>(p->chunk, n_clips - c0));
    const float* x = d_mfcc + c0 * p->H0 * p->W0;
    {
      ProfSpan span("res_conv0", s);
      HONK_CUDA(cudaMemsetAsync(pool, 0, static_cast<size_t>(nb) * p->N * sizeof(float), s));
      if (p->ncg == 6)
        tc_conv0_kernel<6><<<nb, 256, smem0, s>>>(x, p->w0, bufA, p->H0, p->W0, p->pool_h, p->pool_w, p->H, p->W);
      else
        tc_conv0_kernel<4><<<nb, 256, smem0, s>>>(x, p->w0, bufA, p->H0, p->W0, p->pool_h, p->pool_w, p->H, p->W);
      HONK_LAUNCH_CHECK();
    }
    {
      ProfSpan span("res_conv", s);
      for (int i = 1; i <= p->n_layers; ++i) {
        TcConvArgs a;
        a.g = p->geom[i - 1];
        a.n_clips = nb;
        a.flags = ((i % 2 == 0) ? 1 : 0) | ((i == p->n_layers) ? 2 : 0);
        a.wpack = p->wpack[i - 1];
        a.bn_scale = p->bn_scale[i - 1];
        a.bn_shift = p->bn_shift[i - 1];
        a.out = (i % 2 == 1) ? bufB : bufC;
        a.res = bufA;
        a.pool = pool;
        const int grid = std::min(nb * a.g.n_units, p->grid_cap);
        HONK_TRY(tc_conv_launch(p->ncg, a, maps[i - 1], grid, s));
      }
    }
    {
      ProfSpan span("res_tail", s);
      const int total = nb * p->n_labels;
      tc_tail_kernel<<<(total + 127) / 128, 128, 0, s>>>(pool, p->wout, p->bout, d_logits + c0 * p->n_labels, nb, p->N,
                                                         p->n_labels, 1.0f / static_cast<float>(p->H * p->W));
      HONK_LAUNCH_CHECK();
    }
  }
  return HONK_OK;
}

}  // namespace honk
