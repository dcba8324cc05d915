// Convolution entry points (host C++) built on tc_conv / tc_wgrad.
#pragma once
#include "tc_api.h"

namespace act {

struct ConvShape {
  int N, H, W, Cin, Cout, KH, stride, pad;     // square kernel KH x KH; input [N,H,W,Cin] NHWC
};

int conv_out_dim(int in, int k, int stride, int pad);
int conv_fwd(const ConvShape& s, const bf16* x, const bf16* wpk, const EpiDesc& epi, cudaStream_t stream);
int conv_dgrad(const ConvShape& s, const bf16* dy, const bf16* wpkt, EpiDesc epi, const bf16* res_compact,
               cudaStream_t stream);
int conv_wgrad(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
               cudaStream_t stream);

}  // namespace act
