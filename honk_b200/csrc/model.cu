// Model handle: strict state_dict validation (load_state_dict semantics, utils/model.py:79-80),
// the all-FP32 CUDA-core forward for every architecture, and the C ABI.
#include "model.cuh"

#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <vector>

using namespace honk;

namespace {

constexpr int64_t kChunkF32 = 512;   // clips per pass on the FP32 path (bounds the workspace)

struct Expect {
  std::string name;
  std::vector<int64_t> dims;
};

int conv_out(int n, int k, int s) { return (n - k) / s + 1; }

// Key names and shapes of the reference modules (model.py:85-102 and 126-184).
int expected_keys(const honk_model_config& c, std::vector<Expect>& out, std::vector<std::string>& ignorable) {
  if (c.family == HONK_FAMILY_RES) {
    const int C = c.n_feature_maps, L = c.n_labels;
    HONK_CHECK(C >= 1 && C <= 512 && L >= 1 && c.n_layers >= 1 && c.n_layers <= 64, HONK_ERR_INVALID,
               "bad res config (maps=%d labels=%d layers=%d)", C, L, c.n_layers);
    out.push_back({"conv0.weight", {C, 1, 3, 3}});
    for (int i = 1; i <= c.n_layers; ++i) {
      const std::string s = std::to_string(i);
      out.push_back({"bn" + s + ".running_mean", {C}});
      out.push_back({"bn" + s + ".running_var", {C}});
      ignorable.push_back("bn" + s + ".num_batches_tracked");
      out.push_back({"conv" + s + ".weight", {C, C, 3, 3}});
    }
    out.push_back({"output.weight", {L, C}});
    out.push_back({"output.bias", {L}});
    return HONK_OK;
  }
  HONK_CHECK(c.family == HONK_FAMILY_CNN, HONK_ERR_INVALID, "unknown model family %d", c.family);
  HONK_CHECK(c.n_feature_maps1 >= 1 && c.conv1_size[0] >= 1 && c.conv1_size[1] >= 1 && c.conv1_stride[0] >= 1 &&
                 c.conv1_stride[1] >= 1 && c.conv1_pool[0] >= 1 && c.conv1_pool[1] >= 1 && c.n_labels >= 1,
             HONK_ERR_INVALID, "bad cnn config");
  int H = c.height, W = c.width, C = c.n_feature_maps1;
  HONK_CHECK(H >= c.conv1_size[0] && W >= c.conv1_size[1], HONK_ERR_INVALID, "conv1 larger than the input");
  out.push_back({"conv1.weight", {C, 1, c.conv1_size[0], c.conv1_size[1]}});
  out.push_back({"conv1.bias", {C}});
  H = conv_out(H, c.conv1_size[0], c.conv1_stride[0]) / c.conv1_pool[0];
  W = conv_out(W, c.conv1_size[1], c.conv1_stride[1]) / c.conv1_pool[1];
  if (c.has_conv2) {
    HONK_CHECK(c.n_feature_maps2 >= 1 && c.conv2_size[0] >= 1 && c.conv2_size[1] >= 1 && c.conv2_stride[0] >= 1 &&
                   c.conv2_stride[1] >= 1 && c.conv2_pool[0] >= 1 && c.conv2_pool[1] >= 1,
               HONK_ERR_INVALID, "bad conv2 config");
    HONK_CHECK(H >= c.conv2_size[0] && W >= c.conv2_size[1], HONK_ERR_INVALID, "conv2 larger than its input");
    out.push_back({"conv2.weight", {c.n_feature_maps2, C, c.conv2_size[0], c.conv2_size[1]}});
    out.push_back({"conv2.bias", {c.n_feature_maps2}});
    H = conv_out(H, c.conv2_size[0], c.conv2_stride[0]) / c.conv2_pool[0];
    W = conv_out(W, c.conv2_size[1], c.conv2_stride[1]) / c.conv2_pool[1];
    C = c.n_feature_maps2;
  }
  HONK_CHECK(H >= 1 && W >= 1, HONK_ERR_INVALID, "feature map vanished");
  const int64_t flat = static_cast<int64_t>(C) * H * W;
  int64_t last = flat;
  if (!c.tf_variant) {
    out.push_back({"lin.weight", {32, flat}});
    out.push_back({"lin.bias", {32}});
  }
  if (c.dnn1_size > 0) {
    out.push_back({"dnn1.weight", {c.dnn1_size, c.tf_variant ? flat : 32}});
    out.push_back({"dnn1.bias", {c.dnn1_size}});
    last = c.dnn1_size;
    if (c.dnn2_size > 0) {
      out.push_back({"dnn2.weight", {c.dnn2_size, c.dnn1_size}});
      out.push_back({"dnn2.bias", {c.dnn2_size}});
      last = c.dnn2_size;
    }
  }
  out.push_back({"output.weight", {c.n_labels, last}});
  out.push_back({"output.bias", {c.n_labels}});
  return HONK_OK;
}

int upload(honk_model* m, const std::vector<float>& h, float** d) {
  void* p = nullptr;
  HONK_CUDA(cudaMalloc(&p, std::max<size_t>(h.size(), 1) * sizeof(float)));
  m->owned.push_back(p);
  HONK_CUDA(cudaMemcpy(p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  *d = static_cast<float*>(p);
  return HONK_OK;
}

int res_dilation(const honk_model_config& c, int i) {   // conv_i, i >= 1 (model.py:94-98)
  return c.use_dilation ? (1 << ((i - 1) / 3)) : 1;
}

// per-clip float count of the largest activation on the FP32 path
int64_t f32_act_elems(const honk_model* m) {
  const honk_model_config& c = m->cfg;
  if (c.family == HONK_FAMILY_RES) return static_cast<int64_t>(c.n_feature_maps) * m->H0 * m->W0;
  int H = conv_out(c.height, c.conv1_size[0], c.conv1_stride[0]);
  int W = conv_out(c.width, c.conv1_size[1], c.conv1_stride[1]);
  int64_t mx = static_cast<int64_t>(c.n_feature_maps1) * H * W;
  H /= c.conv1_pool[0];
  W /= c.conv1_pool[1];
  if (c.has_conv2) {
    H = conv_out(H, c.conv2_size[0], c.conv2_stride[0]);
    W = conv_out(W, c.conv2_size[1], c.conv2_stride[1]);
    mx = std::max(mx, static_cast<int64_t>(c.n_feature_maps2) * H * W);
  }
  return std::max<int64_t>(mx, 512);
}

int64_t f32_workspace(const honk_model* m, int64_t n) {
  const int64_t chunk = std::min(n, kChunkF32);
  Arena a(nullptr, 0);
  for (int i = 0; i < 4; ++i) a.take<float>(f32_act_elems(m) * chunk);
  return align_up(a.off, 256);
}

int forward_res_f32(const honk_model* m, const float* x, int B, float* logits, Arena& ar, cudaStream_t s) {
  const honk_model_config& c = m->cfg;
  const int C = c.n_feature_maps;
  const int64_t per = f32_act_elems(m) * B;
  float* b0 = ar.take<float>(per);
  float* b1 = ar.take<float>(per);
  float* b2 = ar.take<float>(per);
  float* b3 = ar.take<float>(per);
  ConvGeom g{B, 1, m->H0, m->W0, C, 3, 3, 1, 1, 1, 1, 1, 1, 0, 0};
  conv_out_dims(g);
  float* xin = b1;      // normalised conv input
  float* old = b2;      // raw residual stream
  if (c.res_pool[0] > 0) {
    HONK_TRY(conv2d_f32(x, m->dev.at("conv0.weight"), nullptr, b0, g, true, s));
    HONK_TRY(pool2d_f32(b0, old, B * C, m->H0, m->W0, c.res_pool[0], c.res_pool[1], false, s));
  } else {
    HONK_TRY(conv2d_f32(x, m->dev.at("conv0.weight"), nullptr, old, g, true, s));
  }
  const float* cur = old;   // conv1 reads the un-normalised conv0 output (there is no bn0)
  const int HW = m->Hr * m->Wr;
  for (int i = 1; i <= c.n_layers; ++i) {
    const int d = res_dilation(c, i);
    ConvGeom gi{B, C, m->Hr, m->Wr, C, 3, 3, 1, 1, d, d, d, d, 0, 0};
    conv_out_dims(gi);
    const std::string si = std::to_string(i);
    HONK_TRY(conv2d_f32(cur, m->dev.at("conv" + si + ".weight"), nullptr, b0, gi, true, s));
    float* dst = (cur == xin) ? b3 : xin;
    const bool add = (i % 2 == 0);
    HONK_TRY(bn_residual_f32(b0, add ? old : nullptr, add ? old : nullptr, dst,
                             m->dev.at("bn" + si + ".running_mean"), m->bn_scale[i - 1], B, C, HW, s));
    if (dst == b3) std::swap(xin, b3);
    cur = xin;
  }
  HONK_TRY(mean_hw_f32(cur, b0, B * C, HW, s));
  ConvGeom gl{B, C, 1, 1, c.n_labels, 1, 1, 1, 1, 0, 0, 1, 1, 1, 1};
  return conv2d_f32(b0, m->dev.at("output.weight"), m->dev.at("output.bias"), logits, gl, false, s);
}

int linear_f32(const honk_model* m, const char* name, const float* in, float* out, int B, bool relu, cudaStream_t s) {
  const HostTensor& w = m->host.at(std::string(name) + ".weight");
  ConvGeom g{B, static_cast<int>(w.dims[1]), 1, 1, static_cast<int>(w.dims[0]), 1, 1, 1, 1, 0, 0, 1, 1, 1, 1};
  return conv2d_f32(in, m->dev.at(std::string(name) + ".weight"), m->dev.at(std::string(name) + ".bias"), out, g,
                    relu, s);
}

int forward_cnn_f32(const honk_model* m, const float* x, int B, float* logits, Arena& ar, cudaStream_t s) {
  const honk_model_config& c = m->cfg;
  const int64_t per = f32_act_elems(m) * B;
  float* a = ar.take<float>(per);
  float* b = ar.take<float>(per);
  ConvGeom g{B, 1, c.height, c.width, c.n_feature_maps1, c.conv1_size[0], c.conv1_size[1],
             c.conv1_stride[0], c.conv1_stride[1], 0, 0, 1, 1, 0, 0};
  conv_out_dims(g);
  HONK_TRY(conv2d_f32(x, m->dev.at("conv1.weight"), m->dev.at("conv1.bias"), a, g, true, s));
  int C = g.Cout, H = g.Ho, W = g.Wo;
  float* cur = a;
  float* nxt = b;
  if (c.conv1_pool[0] != 1 || c.conv1_pool[1] != 1) {
    HONK_TRY(pool2d_f32(cur, nxt, B * C, H, W, c.conv1_pool[0], c.conv1_pool[1], true, s));
    std::swap(cur, nxt);
    H /= c.conv1_pool[0];
    W /= c.conv1_pool[1];
  }
  if (c.has_conv2) {
    ConvGeom g2{B, C, H, W, c.n_feature_maps2, c.conv2_size[0], c.conv2_size[1],
                c.conv2_stride[0], c.conv2_stride[1], 0, 0, 1, 1, 0, 0};
    conv_out_dims(g2);
    HONK_TRY(conv2d_f32(cur, m->dev.at("conv2.weight"), m->dev.at("conv2.bias"), nxt, g2, true, s));
    std::swap(cur, nxt);
    C = g2.Cout; H = g2.Ho; W = g2.Wo;
    if (c.conv2_pool[0] != 1 || c.conv2_pool[1] != 1) {
      HONK_TRY(pool2d_f32(cur, nxt, B * C, H, W, c.conv2_pool[0], c.conv2_pool[1], true, s));
      std::swap(cur, nxt);
      H /= c.conv2_pool[0];
      W /= c.conv2_pool[1];
    }
  }
  // flatten is a no-op in NCHW (x.view(B, -1), model.py:195)
  if (!c.tf_variant) {
    HONK_TRY(linear_f32(m, "lin", cur, nxt, B, false, s));
    std::swap(cur, nxt);
  }
  if (c.dnn1_size > 0) {
    HONK_TRY(linear_f32(m, "dnn1", cur, nxt, B, !c.tf_variant, s));
    std::swap(cur, nxt);
    if (c.dnn2_size > 0) {
      HONK_TRY(linear_f32(m, "dnn2", cur, nxt, B, false, s));
      std::swap(cur, nxt);
    }
  }
  return linear_f32(m, "output", cur, logits, B, false, s);
}

}  // namespace

extern "C" int honk_model_create(const honk_model_config* cfg, const honk_tensor* tensors, int32_t n_tensors,
                                 int device, honk_model** out) {
  HONK_CHECK(cfg && out && (tensors || n_tensors == 0), HONK_ERR_INVALID, "honk_model_create: null argument");
  HONK_TRY(honk_device_check(device));
  std::vector<Expect> exp;
  std::vector<std::string> ignorable;
  HONK_TRY(expected_keys(*cfg, exp, ignorable));
  std::unique_ptr<honk_model, int (*)(honk_model*)> m(new honk_model(), honk_model_destroy);
  m->cfg = *cfg;
  m->device = device;
  m->H0 = cfg->height > 0 ? cfg->height : 101;
  m->W0 = cfg->width > 0 ? cfg->width : 40;
  m->cfg.height = m->H0;
  m->cfg.width = m->W0;
  m->Hr = m->H0;
  m->Wr = m->W0;
  if (cfg->family == HONK_FAMILY_RES && cfg->res_pool[0] > 0) {
    HONK_CHECK(cfg->res_pool[1] > 0, HONK_ERR_INVALID, "bad res_pool");
    m->Hr = m->H0 / cfg->res_pool[0];
    m->Wr = m->W0 / cfg->res_pool[1];
    HONK_CHECK(m->Hr >= 1 && m->Wr >= 1, HONK_ERR_INVALID, "res_pool larger than the input");
  }
  // strict key check, like load_state_dict
  for (int i = 0; i < n_tensors; ++i) {
    const honk_tensor& t = tensors[i];
    HONK_CHECK(t.name, HONK_ERR_KEY, "tensor %d has no name", i);
    const std::string name(t.name);
    bool ign = false;
    for (auto& s : ignorable) ign |= (s == name);
    if (ign) continue;
    const Expect* e = nullptr;
    for (auto& x : exp)
      if (x.name == name) e = &x;
    HONK_CHECK(e, HONK_ERR_KEY, "unexpected key \"%s\" in state_dict", t.name);
    HONK_CHECK(!m->host.count(name), HONK_ERR_KEY, "duplicate key \"%s\"", t.name);
    bool same = t.ndim == static_cast<int>(e->dims.size()) && t.h_data != nullptr;
    for (int d = 0; same && d < t.ndim; ++d) same = t.dims[d] == e->dims[d];
    HONK_CHECK(same, HONK_ERR_KEY, "size mismatch for \"%s\"", t.name);
    HostTensor h;
    h.dims = e->dims;
    h.data.assign(static_cast<const float*>(t.h_data), static_cast<const float*>(t.h_data) + h.numel());
    m->host.emplace(name, std::move(h));
  }
  for (auto& x : exp) HONK_CHECK(m->host.count(x.name), HONK_ERR_KEY, "missing key \"%s\" in state_dict", x.name.c_str());

  HONK_CUDA(cudaSetDevice(device));
  m->n_sms = num_sms(device);
  for (auto& kv : m->host) {
    float* d = nullptr;
    HONK_TRY(upload(m.get(), kv.second.data, &d));
    m->dev[kv.first] = d;
  }
  if (cfg->family == HONK_FAMILY_RES) {
    for (int i = 1; i <= cfg->n_layers; ++i) {
      const HostTensor& var = m->host.at("bn" + std::to_string(i) + ".running_var");
      std::vector<float> sc(var.data.size());
      // nn.BatchNorm2d(affine=False) in eval mode, eps = 1e-5 (model.py:100)
      for (size_t k = 0; k < sc.size(); ++k) sc[k] = 1.0f / std::sqrt(var.data[k] + 1e-5f);
      float* d = nullptr;
      HONK_TRY(upload(m.get(), sc, &d));
      m->bn_scale.push_back(d);
    }
    if (cfg->precision == 0) {
      const char* no2 = getenv("HONK_DISABLE_TC2");
      if (!(no2 && atoi(no2)) && cfg->n_feature_maps <= 48) HONK_TRY(res_tc2_plan_create(m.get(), &m->tc2));
      if (!m->tc2) HONK_TRY(res_tc_plan_create(m.get()));
      if (m->tc || m->tc2) m->path = "tcgen05-f16";
    }
  } else if (cfg->precision == 0) {
    static const bool no_ctc = getenv("HONK_DISABLE_CNN_TC") && atoi(getenv("HONK_DISABLE_CNN_TC")) != 0;   // A/B: the per-layer kernels
    if (!no_ctc) HONK_TRY(cnn_tc_plan_create(m.get(), &m->ctc));
    if (!m->ctc) HONK_TRY(cnn_gemm_plan_create(m.get(), &m->cgm));
    if (m->ctc || m->cgm) m->path = "tcgen05-f16";
  }
  *out = m.release();
  return HONK_OK;
}

extern "C" int honk_model_destroy(honk_model* m) {
  if (!m) return HONK_OK;
  if (m->tc) res_tc_plan_free(m->tc);
  if (m->tc2) res_tc2_plan_free(m->tc2);
  if (m->ctc) cnn_tc_plan_free(m->ctc);
  if (m->cgm) cnn_gemm_plan_free(m->cgm);
  for (void* p : m->owned) cudaFree(p);
  delete m;
  return HONK_OK;
}

extern "C" const char* honk_model_path(const honk_model* m) { return m ? m->path.c_str() : ""; }

extern "C" int64_t honk_model_workspace_bytes(const honk_model* m, int64_t n_clips) {
  if (!m || n_clips <= 0) return 0;
  if (m->tc2) return res_tc2_workspace_bytes(m->tc2, n_clips);
  if (m->ctc) return cnn_tc_workspace_bytes(m->ctc, n_clips);
  if (m->cgm) return cnn_gemm_workspace_bytes(m->cgm, n_clips);
  return m->tc ? res_tc_workspace_bytes(m, n_clips) : f32_workspace(m, n_clips);
}

extern "C" int honk_model_forward(const honk_model* m, const float* d_mfcc, int64_t n_clips, int32_t height, int32_t width,
                                  float* d_logits, void* d_workspace, int64_t workspace_bytes, void* stream) {
  HONK_CHECK(m, HONK_ERR_INVALID, "honk_model_forward: null model");
  HONK_CHECK(n_clips >= 0, HONK_ERR_INVALID, "negative n_clips");
  HONK_CHECK(height == m->H0 && width == m->W0, HONK_ERR_INVALID,
             "honk_model_forward: input clips are %d x %d but the model expects %d x %d (frames x coefficients)", height, width,
             m->H0, m->W0);
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(d_mfcc && d_logits, HONK_ERR_INVALID, "honk_model_forward: null buffer");
  HONK_CHECK(d_workspace && workspace_bytes >= honk_model_workspace_bytes(m, n_clips), HONK_ERR_WORKSPACE,
             "workspace too small: %lld < %lld bytes", static_cast<long long>(workspace_bytes),
             static_cast<long long>(honk_model_workspace_bytes(m, n_clips)));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (m->tc2) return res_tc2_forward(m->tc2, d_mfcc, n_clips, d_logits, d_workspace, workspace_bytes, s);
  if (m->ctc) return cnn_tc_forward(m->ctc, d_mfcc, n_clips, d_logits, d_workspace, workspace_bytes, s);
  if (m->cgm) return cnn_gemm_forward(m->cgm, d_mfcc, n_clips, d_logits, d_workspace, workspace_bytes, s);
  if (m->tc) return res_tc_forward(m, d_mfcc, n_clips, d_logits, d_workspace, workspace_bytes, s);
  ProfSpan span("f32_model", s);
  const int64_t in_per = static_cast<int64_t>(m->H0) * m->W0;
  for (int64_t c0 = 0; c0 < n_clips; c0 += kChunkF32) {
    const int B = static_cast<int>(std::min(kChunkF32, n_clips - c0));
    Arena ar(d_workspace, workspace_bytes);
    const float* x = d_mfcc + c0 * in_per;
    float* y = d_logits + c0 * m->cfg.n_labels;
    if (m->cfg.family == HONK_FAMILY_RES) HONK_TRY(forward_res_f32(m, x, B, y, ar, s));
    else HONK_TRY(forward_cnn_f32(m, x, B, y, ar, s));
    HONK_CHECK(ar.ok(), HONK_ERR_WORKSPACE, "internal: workspace overrun");
  }
  return HONK_OK;
}

extern "C" int64_t honk_model_workspace_bytes_pcm(const honk_model* m, const honk_frontend* fe, int64_t n_clips,
                                                  int64_t n_samples) {
  if (!m || !fe || n_clips <= 0) return 0;
  const int64_t mf = align_up(n_clips * honk_frontend_num_frames(fe, n_samples) * m->W0 * 4, 256);
  const int64_t few = honk_frontend_workspace_bytes(fe, n_clips, n_samples, HONK_PCM_I16), mw = honk_model_workspace_bytes(m, n_clips);
  // two layouts: [MFCC][model scratch, shared with the front end's scratch (dead once the MFCCs exist)], or, when the front
  // end writes conv0's operand image straight into the model's scratch, [front-end scratch][model scratch]
  return std::max(mf + std::max(mw, few), few + mw);
}

extern "C" int honk_model_forward_pcm(const honk_model* m, const honk_frontend* fe, const void* d_pcm, int dtype,
                                      int64_t n_clips, int64_t n_samples, int64_t clip_stride, float* d_logits,
                                      void* d_workspace, int64_t workspace_bytes, void* stream) {
  HONK_CHECK(m && fe, HONK_ERR_INVALID, "honk_model_forward_pcm: null handle");
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(honk_frontend_num_frames(fe, n_samples) == m->H0, HONK_ERR_INVALID,
             "front end yields %lld frames but the model expects %d",
             static_cast<long long>(honk_frontend_num_frames(fe, n_samples)), m->H0);
  HONK_CHECK(fe_n_dct(fe) == m->W0, HONK_ERR_INVALID, "front end yields %d coefficients but the model expects %d",
             fe_n_dct(fe), m->W0);
  HONK_CHECK(d_workspace && workspace_bytes >= honk_model_workspace_bytes_pcm(m, fe, n_clips, n_samples),
             HONK_ERR_WORKSPACE, "workspace too small");
  if (m->tc2 && fe_can_x0(fe, dtype, n_clips, n_samples) && !getenv("HONK_NO_X0_FUSION")) {
    // un-pooled res* on the persistent kernel, int16 PCM: the front end's last stage writes conv0's f16 hi/lo operand image
    // itself - no MFCC tensor, no pack kernel (manage_audio.py:30-42 -> model.py:107-111 without the HBM round trip)
    const int64_t few = honk_frontend_workspace_bytes(fe, n_clips, n_samples, dtype);
    char* mws = static_cast<char*>(d_workspace) + few;
    int H = 0, W = 0;
    if (unsigned char* x0 = res_tc2_x0_buffer(m->tc2, n_clips, mws, workspace_bytes - few, &H, &W)) {
      HONK_TRY(fe_forward_x0(fe, d_pcm, dtype, n_clips, n_samples, clip_stride, x0, H, W, d_workspace, few, stream));
      return res_tc2_forward(m->tc2, nullptr, n_clips, d_logits, mws, workspace_bytes - few, static_cast<cudaStream_t>(stream));
    }
  }
  const int64_t mf = align_up(n_clips * static_cast<int64_t>(m->H0) * m->W0 * 4, 256);
  float* mfcc = static_cast<float*>(d_workspace);
  HONK_TRY(honk_frontend_mfcc(fe, d_pcm, dtype, n_clips, n_samples, clip_stride, mfcc, static_cast<char*>(d_workspace) + mf,
                              workspace_bytes - mf, stream));
  return honk_model_forward(m, mfcc, n_clips, m->H0, m->W0, d_logits, static_cast<char*>(d_workspace) + mf,
                            workspace_bytes - mf, stream);
}

// ------------------------------------------------------------------------------------------
// /listen: sliding windows over one buffer (server.py:99-112, service.py:106-110)
// ------------------------------------------------------------------------------------------
namespace {

// One thread per label sums the max-probabilities of the windows it won, in window order and in double like
// the Python loop's float accumulation (server.py:105-108); thread 0 also finds the first window that would
// have ended the command_tagging loop (server.py:109-110).
__global__ void listen_reduce_kernel(const int* __restrict__ labels, const float* __restrict__ probs, long long n_windows,
                                     int n_labels, int keyword, float min_prob, double* __restrict__ sums,
                                     int* __restrict__ counts, int* __restrict__ first_hit) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l < n_labels) {
    double acc = 0.0;
    int cnt = 0;
    for (long long w = 0; w < n_windows; ++w)
      if (labels[w] == l) {
        acc += static_cast<double>(probs[w]);
        ++cnt;
      }
    sums[l] = acc;
    counts[l] = cnt;
  }
  if (l == 0) {
    int hit = -1;
    for (long long w = 0; w < n_windows && hit < 0; ++w)
      if (labels[w] == keyword && probs[w] >= min_prob) hit = static_cast<int>(w);
    *first_hit = hit;
  }
}

int64_t listen_windows(int64_t n_total, int64_t window, int64_t hop) {
  return (window > 0 && hop > 0 && n_total >= window) ? 1 + (n_total - window) / hop : 0;
}

}  // namespace

extern "C" int64_t honk_listen_num_windows(int64_t n_total_samples, int64_t window, int64_t hop) {
  return listen_windows(n_total_samples, window, hop);
}

extern "C" int64_t honk_listen_workspace_bytes(const honk_model* m, const honk_frontend* fe, int64_t n_total_samples,
                                               int64_t window, int64_t hop) {
  const int64_t nw = listen_windows(n_total_samples, window, hop);
  if (!m || !fe || nw <= 0) return 0;
  return align_up(nw * m->cfg.n_labels * 4, 256) + 2 * align_up(nw * 4, 256) + honk_model_workspace_bytes_pcm(m, fe, nw, window);
}

extern "C" int honk_listen(const honk_model* m, const honk_frontend* fe, const void* d_pcm, int dtype, int64_t n_total_samples,
                           int64_t window, int64_t hop, int32_t keyword, float min_keyword_prob, double* d_sums,
                           int32_t* d_counts, int32_t* d_first_hit, void* d_workspace, int64_t workspace_bytes, void* stream) {
  HONK_CHECK(m && fe, HONK_ERR_INVALID, "honk_listen: null handle");
  HONK_CHECK(window > 0 && hop > 0 && n_total_samples >= 0, HONK_ERR_INVALID, "honk_listen: bad window/hop");
  HONK_CHECK(d_sums && d_counts && d_first_hit, HONK_ERR_INVALID, "honk_listen: null output");
  const int64_t nw = listen_windows(n_total_samples, window, hop);
  const int L = m->cfg.n_labels;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nw == 0) {                                  // shorter than one window: the reference loop runs zero times
    HONK_CUDA(cudaMemsetAsync(d_sums, 0, sizeof(double) * L, s));
    HONK_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * L, s));
    HONK_CUDA(cudaMemsetAsync(d_first_hit, 0xFF, sizeof(int32_t), s));
    return HONK_OK;
  }
  HONK_CHECK(d_pcm, HONK_ERR_INVALID, "honk_listen: null buffer");
  HONK_CHECK(d_workspace && workspace_bytes >= honk_listen_workspace_bytes(m, fe, n_total_samples, window, hop),
             HONK_ERR_WORKSPACE, "workspace too small");
  Arena ar(d_workspace, workspace_bytes);
  float* logits = ar.take<float>(nw * L);
  int* labels = ar.take<int>(nw);
  float* probs = ar.take<float>(nw);
  const int64_t used = align_up(ar.off, 256);
  char* inner = static_cast<char*>(d_workspace) + used;
  // window w = samples [w*hop, w*hop + window): rows of the same buffer with clip_stride = hop
  HONK_TRY(honk_model_forward_pcm(m, fe, d_pcm, dtype, nw, window, hop, logits, inner, workspace_bytes - used, stream));
  HONK_TRY(honk_label_from_logits(logits, nw, L, labels, probs, stream));
  listen_reduce_kernel<<<(L + 127) / 128, 128, 0, s>>>(labels, probs, nw, L, keyword, min_keyword_prob, d_sums, d_counts,
                                                       d_first_hit);
  HONK_LAUNCH_CHECK();
  return HONK_OK;
}

// ------------------------------------------------------------------------------------------
// host-buffer pipeline
// ------------------------------------------------------------------------------------------
namespace {

struct HostPipe {                       // copy stream + events of the host pipeline; one per calling thread and device, kept
  cudaStream_t copy = nullptr;          // across calls (creating and destroying them costs ~0.1 ms per call) and released
  std::vector<cudaEvent_t> copied;      // when the thread exits.  A call synchronises both streams before it returns, so
  cudaEvent_t start = nullptr;          // the next call of the same thread can reuse them.
  int device = -1;
  int init(size_t n_chunks) {
    int dev = 0;
    HONK_CUDA(cudaGetDevice(&dev));
    if (dev != device) {
      release();
      device = dev;
    }
    if (!copy) HONK_CUDA(cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking));
    if (!start) HONK_CUDA(cudaEventCreateWithFlags(&start, cudaEventDisableTiming));
    while (copied.size() < n_chunks) {
      cudaEvent_t e = nullptr;
      HONK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      copied.push_back(e);
    }
    return HONK_OK;
  }
  void release() {
    for (cudaEvent_t e : copied)
      if (e) cudaEventDestroy(e);
    copied.clear();
    if (start) cudaEventDestroy(start);
    if (copy) cudaStreamDestroy(copy);
    start = nullptr;
    copy = nullptr;
  }
  ~HostPipe() { release(); }
};

int pcm_elem_bytes(int dtype) { return dtype == HONK_PCM_I16 ? 2 : (dtype == HONK_PCM_F32 ? 4 : 8); }

// Chunk schedule of the host pipeline.  The whole batch is staged in device memory, so the copy stream never waits
// for a buffer; PCIe moves a clip about four times faster than the kernels consume it, so only the FIRST chunk's copy
// is exposed.  It is therefore small (HONK_HOST_FIRST_WAVES, default 4 waves) and the following chunks grow by 3x
// (each one's copy still ends before the previous one's kernels do), up to half the batch.  Every chunk is a whole
// number of waves - the clip count at which all teams of the persistent conv kernel are busy - so chunking costs no
// partially filled wave.  HONK_HOST_CHUNKS=n forces n equal chunks instead.
// Rough device time per clip in microseconds (front end 0.3 + convolutions at ~700 algorithmic TFLOP/s on the
// tcgen05 paths), only used to tell copy-bound models from compute-bound ones.
double est_device_us_per_clip(const honk_model* m) {
  const honk_model_config& c = m->cfg;
  if (m->tc2 || m->tc) return 0.3 + 18.0 * c.n_feature_maps * c.n_feature_maps * m->Hr * m->Wr * c.n_layers / 7e8;
  if (m->ctc || m->cgm) return 0.6;
  return 20.0;
}

std::vector<int64_t> host_chunk_plan(const honk_model* m, int64_t n_clips, int64_t clip_bytes = 32000) {
  std::vector<int64_t> plan;
  if (n_clips <= 0) return plan;
  const int64_t wave = m->tc2 ? honk::res_tc2_wave_clips(m->tc2) : (m->ctc ? m->n_sms : 128);   // cnn_tc_kernel: one clip per CTA at a time
  if (!getenv("HONK_HOST_CHUNKS") && !getenv("HONK_HOST_FIRST_WAVES") &&
      est_device_us_per_clip(m) < 1.5 * static_cast<double>(clip_bytes) / 50e3) {
    // Copy-bound (res8, cnn-trad-pool2: PCIe at ~50 GB/s needs longer per clip than the kernels): the call ends one
    // chunk's kernels after the last copy, so the chunks are equal and small - eight, each a whole number of waves.
    const int64_t chunk = std::max(wave, (n_clips / 8 + wave - 1) / wave * wave);
    for (int64_t c0 = 0; c0 < n_clips; c0 += chunk) plan.push_back(std::min(chunk, n_clips - c0));
    if (plan.size() > 1 && plan.back() < wave / 2) {         // no sliver at the end
      const int64_t last = plan.back();
      plan.pop_back();
      plan.back() += last;
    }
    return plan;
  }
  if (const char* e = getenv("HONK_HOST_CHUNKS")) {
    const int pieces = std::max(1, std::min(64, atoi(e)));
    const int64_t chunk = std::min(n_clips, ((n_clips + pieces - 1) / pieces + wave - 1) / wave * wave);
    for (int64_t c0 = 0; c0 < n_clips; c0 += chunk) plan.push_back(std::min(chunk, n_clips - c0));
    return plan;
  }
  int64_t first = 4;
  if (const char* e = getenv("HONK_HOST_FIRST_WAVES")) first = std::max(1, std::min(1 << 20, atoi(e)));
  int64_t growth = 3;
  if (const char* e = getenv("HONK_HOST_GROWTH")) growth = std::max(2, std::min(64, atoi(e)));
  int64_t cap_pct = 50;
  if (const char* e = getenv("HONK_HOST_CAP_PCT")) cap_pct = std::max(10, std::min(100, atoi(e)));
  const int64_t cap = std::max(wave, (n_clips * cap_pct / 100 + wave - 1) / wave * wave);
  int64_t next = first * wave, left = n_clips;
  while (left > 0) {
    int64_t c = std::min(std::min(next, cap), left);
    if (left - c < first * wave) c = left;                 // no sliver at the end
    plan.push_back(c);
    left -= c;
    next *= growth;
  }
  return plan;
}

}  // namespace

extern "C" int64_t honk_model_workspace_bytes_pcm_host(const honk_model* m, const honk_frontend* fe, int64_t n_clips,
                                                       int64_t n_samples, int dtype) {
  if (!m || !fe || n_clips <= 0) return 0;
  const std::vector<int64_t> plan = host_chunk_plan(m, n_clips, n_samples * pcm_elem_bytes(dtype));
  const int64_t largest = *std::max_element(plan.begin(), plan.end());
  const int64_t staging = align_up(n_clips * n_samples * pcm_elem_bytes(dtype), 256);
  const int64_t logits = align_up(n_clips * m->cfg.n_labels * 4, 256);
  return staging + logits + 2 * align_up(n_clips * 4, 256) + honk_model_workspace_bytes_pcm(m, fe, largest, n_samples);
}

namespace {
// host PCM -> host logits and / or host (label, prob): the body of both host-buffer entry points
int forward_pcm_host_impl(const honk_model* m, const honk_frontend* fe, const void* h_pcm, int dtype, int64_t n_clips,
                          int64_t n_samples, int64_t clip_stride, float* h_logits, int32_t* h_labels, float* h_probs,
                          void* d_workspace, int64_t workspace_bytes, void* stream) {
  HONK_CHECK(m && fe, HONK_ERR_INVALID, "honk_model_forward_pcm_host: null handle");
  HONK_CHECK(n_clips >= 0 && n_samples > 0 && clip_stride >= n_samples, HONK_ERR_INVALID,
             "honk_model_forward_pcm_host: bad shape");
  if (n_clips == 0) return HONK_OK;
  HONK_CHECK(h_pcm && (h_logits || (h_labels && h_probs)), HONK_ERR_INVALID, "honk_model_forward_pcm_host: null buffer");
  HONK_CHECK(dtype == HONK_PCM_I16 || dtype == HONK_PCM_F32 || dtype == HONK_PCM_F64, HONK_ERR_INVALID, "unknown PCM dtype %d", dtype);
  HONK_CHECK(d_workspace && workspace_bytes >= honk_model_workspace_bytes_pcm_host(m, fe, n_clips, n_samples, dtype),
             HONK_ERR_WORKSPACE, "workspace too small");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int eb = pcm_elem_bytes(dtype);
  const std::vector<int64_t> plan = host_chunk_plan(m, n_clips, n_samples * eb);
  const int64_t staging = align_up(n_clips * n_samples * eb, 256);
  const int64_t logits_bytes = align_up(n_clips * m->cfg.n_labels * 4, 256);
  char* base = static_cast<char*>(d_workspace);
  float* d_logits = reinterpret_cast<float*>(base + staging);
  const int64_t lab_bytes = align_up(n_clips * 4, 256);
  int32_t* d_labels = reinterpret_cast<int32_t*>(base + staging + logits_bytes);
  float* d_probs = reinterpret_cast<float*>(base + staging + logits_bytes + lab_bytes);
  char* inner = base + staging + logits_bytes + 2 * lab_bytes;
  const int64_t inner_bytes = workspace_bytes - staging - logits_bytes - 2 * lab_bytes;
  auto finish = [&]() -> int {
    if (h_logits)
      HONK_CUDA(cudaMemcpyAsync(h_logits, d_logits, static_cast<size_t>(n_clips * m->cfg.n_labels * 4), cudaMemcpyDeviceToHost, s));
    if (h_labels && h_probs) {                               // labels and probs are adjacent: one copy when the host side is too
      HONK_TRY(honk_label_from_logits(d_logits, n_clips, m->cfg.n_labels, d_labels, d_probs, stream));
      HONK_CUDA(cudaMemcpyAsync(h_labels, d_labels, static_cast<size_t>(n_clips * 4), cudaMemcpyDeviceToHost, s));
      HONK_CUDA(cudaMemcpyAsync(h_probs, d_probs, static_cast<size_t>(n_clips * 4), cudaMemcpyDeviceToHost, s));
    }
    HONK_CUDA(cudaStreamSynchronize(s));
    return HONK_OK;
  };
  if (plan.size() == 1) {
    // a request-sized batch: one copy and the kernels on the caller's stream, no second stream or events
    if (clip_stride == n_samples) {
      HONK_CUDA(cudaMemcpyAsync(base, h_pcm, static_cast<size_t>(n_clips * n_samples * eb), cudaMemcpyHostToDevice, s));
    } else {
      HONK_CUDA(cudaMemcpy2DAsync(base, static_cast<size_t>(n_samples * eb), h_pcm, static_cast<size_t>(clip_stride * eb),
                                  static_cast<size_t>(n_samples * eb), static_cast<size_t>(n_clips), cudaMemcpyHostToDevice, s));
    }
    HONK_TRY(honk_model_forward_pcm(m, fe, base, dtype, n_clips, n_samples, n_samples, d_logits, inner, inner_bytes, s));
    return finish();
  }
  // everything that can be refused is checked before the first copy is queued
  HONK_CHECK(honk_frontend_num_frames(fe, n_samples) == m->H0 && fe_n_dct(fe) == m->W0, HONK_ERR_INVALID,
             "front end yields %lld x %d but the model expects %d x %d", static_cast<long long>(honk_frontend_num_frames(fe, n_samples)),
             fe_n_dct(fe), m->H0, m->W0);
  static thread_local HostPipe hp;
  HONK_TRY(hp.init(plan.size()));
  // From here on copies into the caller's workspace are in flight on the copy stream: a failure below must not return
  // while they run (the caller may free or reuse the workspace and the pinned buffers), and the thread's events must be
  // quiescent for its next call.  The guard drains both streams on every early return.
  struct Drain {
    cudaStream_t a, b;
    bool armed = true;
    ~Drain() {
      if (armed) {
        cudaStreamSynchronize(a);
        cudaStreamSynchronize(b);
      }
    }
  } drain{hp.copy, s};
  // the scratch may have been handed out in the order of `s`: the copy stream must not run ahead of that
  HONK_CUDA(cudaEventRecord(hp.start, s));
  HONK_CUDA(cudaStreamWaitEvent(hp.copy, hp.start, 0));
  int64_t c0 = 0;
  for (size_t i = 0; i < plan.size(); ++i) {                // all copies are queued first: the copy engine runs ahead
    const int64_t nc = plan[i];
    const char* src = static_cast<const char*>(h_pcm) + c0 * clip_stride * eb;
    char* dst = base + c0 * n_samples * eb;
    if (clip_stride == n_samples) {
      HONK_CUDA(cudaMemcpyAsync(dst, src, static_cast<size_t>(nc * n_samples * eb), cudaMemcpyHostToDevice, hp.copy));
    } else {
      HONK_CUDA(cudaMemcpy2DAsync(dst, static_cast<size_t>(n_samples * eb), src, static_cast<size_t>(clip_stride * eb),
                                  static_cast<size_t>(n_samples * eb), static_cast<size_t>(nc), cudaMemcpyHostToDevice, hp.copy));
    }
    HONK_CUDA(cudaEventRecord(hp.copied[i], hp.copy));
    c0 += nc;
  }
  c0 = 0;
  for (size_t i = 0; i < plan.size(); ++i) {
    const int64_t nc = plan[i];
    HONK_CUDA(cudaStreamWaitEvent(s, hp.copied[i], 0));
    HONK_TRY(honk_model_forward_pcm(m, fe, base + c0 * n_samples * eb, dtype, nc, n_samples, n_samples,
                                    d_logits + c0 * m->cfg.n_labels, inner, inner_bytes, s));
    c0 += nc;
  }
  HONK_TRY(finish());
  HONK_CUDA(cudaStreamSynchronize(hp.copy));
  drain.armed = false;                                       // both streams are already idle
  return HONK_OK;
}
}  // namespace

extern "C" int honk_model_forward_pcm_host(const honk_model* m, const honk_frontend* fe, const void* h_pcm, int dtype,
                                           int64_t n_clips, int64_t n_samples, int64_t clip_stride, float* h_logits,
                                           void* d_workspace, int64_t workspace_bytes, void* stream) {
  if (n_clips > 0) HONK_CHECK(h_logits, HONK_ERR_INVALID, "honk_model_forward_pcm_host: null buffer");
  return forward_pcm_host_impl(m, fe, h_pcm, dtype, n_clips, n_samples, clip_stride, h_logits, nullptr, nullptr, d_workspace,
                               workspace_bytes, stream);
}

extern "C" int honk_model_label_pcm_host(const honk_model* m, const honk_frontend* fe, const void* h_pcm, int dtype,
                                         int64_t n_clips, int64_t n_samples, int64_t clip_stride, int32_t* h_labels,
                                         float* h_probs, void* d_workspace, int64_t workspace_bytes, void* stream) {
  if (n_clips > 0) HONK_CHECK(h_labels && h_probs, HONK_ERR_INVALID, "honk_model_label_pcm_host: null buffer");
  return forward_pcm_host_impl(m, fe, h_pcm, dtype, n_clips, n_samples, clip_stride, nullptr, h_labels, h_probs, d_workspace,
                               workspace_bytes, stream);
}
