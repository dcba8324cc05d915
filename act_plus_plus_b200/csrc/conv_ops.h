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
int conv_wgrad_reserve(long long elems);     // staging buffer [KK][Cout][Cin] fp32 (allocate before any stream capture)
struct FoldJob { const float* T; const float* scale; float* dw; float alpha; int Cout, Cin, KK, tile0; };
struct FoldJobs { static constexpr int MAX = 64; int n = 0; FoldJob job[MAX]; };
int conv_wgrad_staged(const ConvShape& s, const bf16* x, const bf16* dy, float* T, cudaStream_t stream);
int wgrad_fold_batched(FoldJobs& jobs, cudaStream_t stream);   // dW[co][ci][kk] += alpha*scale[co]*T[kk][co][ci], all jobs in one launch
int conv_wgrad(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
               cudaStream_t stream);

}  // namespace act
