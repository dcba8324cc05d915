// Convolution entry points (host C++) built on tc_conv / tc_wgrad.
#pragma once
#include "tc_api.h"

namespace act {

struct ConvShape {
  int N, H, W, Cin, Cout, KH, stride, pad;     // square kernel KH x KH; input [N,H,W,Cin] NHWC
};

int conv_out_dim(int in, int k, int stride, int pad);
// wpair (optional, 64 -> 64 channel 3x3/s1 convs with an even width): the paired operand [128][9*128] built by
// pack_pair_batched with conv_pair_slots(); one GEMM row then produces two horizontally adjacent output pixels (N = 128
// at the MMA count of 1.5 N=64 rows).  Per-channel epilogue vectors (scale / bias) must then hold 128 entries (64 + 64).
int conv_fwd(const ConvShape& s, const bf16* x, const bf16* wpk, const EpiDesc& epi, cudaStream_t stream,
             const bf16* wpair = nullptr);
int conv_dgrad(const ConvShape& s, const bf16* dy, const bf16* wpkt, EpiDesc epi, const bf16* res_compact,
               cudaStream_t stream, const bf16* wpair = nullptr);
bool conv_pair_ok(const ConvShape& s);
// source slots (kh*3+kw of wpk for fwd, of wpkt for dgrad) of the 18 paired blocks (tap*2 + input pixel parity), for the even / odd pixel rows
void conv_pair_slots(bool dgrad, int8_t even_slot[18], int8_t odd_slot[18]);
int conv_wgrad_reserve(long long elems);     // staging buffer [KK][Cout][Cin] fp32 (allocate before any stream capture)
struct FoldJob { const float* T; const float* scale; float* dw; float alpha; int Cout, Cin, KK, tile0; };
struct FoldJobs { static constexpr int MAX = 64; int n = 0; FoldJob job[MAX]; };
int conv_wgrad_staged(const ConvShape& s, const bf16* x, const bf16* dy, float* T, cudaStream_t stream);
int wgrad_fold_batched(FoldJobs& jobs, cudaStream_t stream);   // dW[co][ci][kk] += alpha*scale[co]*T[kk][co][ci], all jobs in one launch
// allocation-free variant (static device scratch, KK*Cin*Cout <= CONV_WGRAD_NOALLOC_MAX; calls overlapping on the GPU race)
constexpr long long CONV_WGRAD_NOALLOC_MAX = 9LL * 256 * 256;
int conv_wgrad_noalloc(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
                       cudaStream_t stream);
int conv_wgrad(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
               cudaStream_t stream);

}  // namespace act
