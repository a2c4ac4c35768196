// Model handle shared by the FP32 CUDA-core path (model.cu) and the tcgen05 path (res_tc.cu).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "generic_ops.cuh"

namespace honk {

struct HostTensor {
  std::vector<int64_t> dims;
  std::vector<float> data;
  int64_t numel() const {
    int64_t n = 1;
    for (auto d : dims) n *= d;
    return n;
  }
};

struct ResTcPlan;   // res_tc.cu
void res_tc_plan_free(ResTcPlan*);
struct ResTc2Plan;  // res_tc2.cu
void res_tc2_plan_free(ResTc2Plan*);
struct CnnTcPlan;   // cnn_tc.cu
void cnn_tc_plan_free(CnnTcPlan*);
struct CnnGemmPlan; // gemm_tc.cu
void cnn_gemm_plan_free(CnnGemmPlan*);

}  // namespace honk

struct honk_model {
  honk_model_config cfg{};
  int device = 0;
  int n_sms = 148;
  std::map<std::string, honk::HostTensor> host;      // validated state_dict (float32)
  std::map<std::string, float*> dev;                 // same tensors on the device (FP32 path)
  std::vector<float*> bn_scale;                      // [n_layers] 1/sqrt(var + 1e-5), device
  std::vector<void*> owned;                          // every cudaMalloc of this handle
  honk::ResTcPlan* tc = nullptr;                     // non-null => per-layer tcgen05 path (pooled res8/res26)
  honk::ResTc2Plan* tc2 = nullptr;                   // non-null => persistent multi-layer tcgen05 path (res15)
  honk::CnnTcPlan* ctc = nullptr;                    // non-null => tcgen05 path for the cnn-trad-pool2 geometry
  honk::CnnGemmPlan* cgm = nullptr;                  // non-null => generic tcgen05 GEMM path for the other cnn-* configs
  std::string path = "cuda-core-f32";
  // geometry derived from cfg
  int H0 = 101, W0 = 40;                             // input
  int Hr = 101, Wr = 40;                             // res*: map size after the optional pool
};

namespace honk {

// res_tc.cu: builds the tensor-core plan when the architecture has one (res* models).
// Returns HONK_OK with m->tc == nullptr when it does not apply.
int fe_n_dct(const honk_frontend* fe);   // frontend.cu
int res_tc_plan_create(honk_model* m);
int64_t res_tc_workspace_bytes(const honk_model* m, int64_t n_clips);
// res_tc2.cu: persistent multi-layer kernel for the un-pooled res* models; *out == nullptr when it does not apply.
int res_tc2_plan_create(honk_model* m, ResTc2Plan** out);
int64_t res_tc2_workspace_bytes(const ResTc2Plan* p, int64_t n_clips);
int res_tc2_wave_clips(const ResTc2Plan* p);   // clips all teams process at once: batch cuts on this boundary lose nothing
// Un-pooled models: where conv0's operand image x0[clip][W][H + 2][8 halves] of this batch lives inside `ws` (so that the
// front end can write it directly); nullptr for pooled models.  d_mfcc == nullptr in res_tc2_forward: x0 is already there.
unsigned char* res_tc2_x0_buffer(const ResTc2Plan* p, int64_t n_clips, void* ws, int64_t ws_bytes, int* H, int* W);
int fe_forward_x0(const honk_frontend* fe, const void* d_pcm, int dtype, int64_t n_clips, int64_t n_samples, int64_t clip_stride,
                  unsigned char* d_x0, int H, int W, void* d_workspace, int64_t workspace_bytes, void* stream);   // frontend.cu
bool fe_can_x0(const honk_frontend* fe, int dtype, int64_t n_clips, int64_t n_samples);
int res_tc2_forward(const ResTc2Plan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                    cudaStream_t s);
// gemm_tc.cu: every layer of a cnn-* model as one tcgen05 GEMM (im2col pack + gemm_tc_kernel)
int cnn_gemm_plan_create(honk_model* m, CnnGemmPlan** out);
int64_t cnn_gemm_workspace_bytes(const CnnGemmPlan* p, int64_t n_clips);
int cnn_gemm_forward(const CnnGemmPlan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                     cudaStream_t s);
// cnn_tc.cu: tcgen05 conv1 + conv2 for cnn-trad-pool2; *out == nullptr when the geometry differs.
int cnn_tc_plan_create(honk_model* m, CnnTcPlan** out);
int64_t cnn_tc_workspace_bytes(const CnnTcPlan* p, int64_t n_clips);
int cnn_tc_forward(const CnnTcPlan* p, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws, int64_t ws_bytes,
                   cudaStream_t s);
int res_tc_forward(const honk_model* m, const float* d_mfcc, int64_t n_clips, float* d_logits, void* ws,
                   int64_t ws_bytes, cudaStream_t s);

}  // namespace honk
