// Host-callable launchers of the bandwidth-bound kernels (pack / normalise / softmax / loss / optimizer ...).
#pragma once
#include <algorithm>
#include "tc_api.h"

namespace act {

int cast_bf16(const float* src, bf16* dst, long long n, cudaStream_t st);
int pack_conv_weights(const float* w, const float* scale, bf16* wpk, bf16* wpkt, int Cout, int Cin, int KK, cudaStream_t st);
int bn_fold(const float* w, const float* b, const float* rm, const float* rv, float* scale, float* shift, int n, cudaStream_t st);

}  // namespace act
