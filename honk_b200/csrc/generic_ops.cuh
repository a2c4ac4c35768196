// All-FP32 CUDA-core operators (NCHW).  They back every architecture of _configs
// (utils/model.py:381-414) on the "cuda-core-f32" path and are the on-GPU numerical reference
// for the tcgen05 path.  Host launchers return a honk_status.
#pragma once
#include "common.cuh"

namespace honk {

struct ConvGeom {
  int B, Cin, H, W;          // input  [B,Cin,H,W]
  int Cout, KH, KW;          // weight [Cout,Cin,KH,KW]
  int sh, sw, ph, pw, dh, dw;
  int Ho, Wo;                // output [B,Cout,Ho,Wo]
};

inline void conv_out_dims(ConvGeom& g) {
  g.Ho = (g.H + 2 * g.ph - g.dh * (g.KH - 1) - 1) / g.sh + 1;
  g.Wo = (g.W + 2 * g.pw - g.dw * (g.KW - 1) - 1) / g.sw + 1;
}

// out = act(conv2d(in, w) + bias); a Linear layer is the KH=KW=H=W=1 case.
int conv2d_f32(const float* in, const float* w, const float* bias, float* out, const ConvGeom& g, bool relu,
               cudaStream_t s);
// kernel == stride, floor mode (nn.MaxPool2d / nn.AvgPool2d defaults)
int pool2d_f32(const float* in, float* out, int BC, int H, int W, int kh, int kw, bool is_max, cudaStream_t s);
// x = y (+ res if res != nullptr); if (res_out) res_out = x; out = (x - mean[c]) * scale[c]
int bn_residual_f32(const float* y, const float* res, float* res_out, float* out, const float* mean,
                    const float* scale, int B, int C, int HW, cudaStream_t s);
// out[b,c] = mean over HW
int mean_hw_f32(const float* in, float* out, int BC, int HW, cudaStream_t s);

}  // namespace honk
