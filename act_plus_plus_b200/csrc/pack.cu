// Weight shadow kernels: fp32 master parameters -> bf16 operands in kernel-native layouts.
#include "kernels.h"

namespace act {

__global__ void cast_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, long long n) {
  long long i = (blockIdx.x * 1LL * blockDim.x + threadIdx.x) * 4;
  const long long stride = gridDim.x * 1LL * blockDim.x * 4;
  for (; i + 3 < n; i += stride) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&a); o.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(dst + i) = o;
  }
  if (i < n) for (long long j = i; j < n && j < i + 4; ++j) dst[j] = __float2bfloat16_rn(src[j]);
}

int cast_bf16(const float* src, bf16* dst, long long n, cudaStream_t st) {
  if (n <= 0) return 0;
  const int threads = 256;
  const long long want = (n / 4 + threads - 1) / threads + 1;
  const int blocks = static_cast<int>(want < 148 * 8 ? want : 148 * 8);
  cast_bf16_kernel<<<blocks, threads, 0, st>>>(src, dst, n);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// w [Cout][Cin][KK] fp32 -> wpk [Cout][KK*Cin] (tap-major then ci) and wpkt [Cin][KK*Cout] * scale[co]
__global__ void pack_conv_kernel(const float* __restrict__ w, const float* __restrict__ scale, bf16* __restrict__ wpk,
                                 bf16* __restrict__ wpkt, int Cout, int Cin, int KK) {
  const long long total = 1LL * Cout * Cin * KK;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    // iterate in wpk order so the bf16 stores coalesce: i = (co*KK + t)*Cin + ci
    const int ci = static_cast<int>(i % Cin);
    const long long r = i / Cin;
    const int t = static_cast<int>(r % KK), co = static_cast<int>(r / KK);
    const float v = w[(1LL * co * Cin + ci) * KK + t];
    if (wpk) wpk[i] = __float2bfloat16_rn(v);
    if (wpkt) wpkt[(1LL * ci * KK + t) * Cout + co] = __float2bfloat16_rn(scale ? v * scale[co] : v);
  }
}

int pack_conv_weights(const float* w, const float* scale, bf16* wpk, bf16* wpkt, int Cout, int Cin, int KK, cudaStream_t st) {
  const long long total = 1LL * Cout * Cin * KK;
  const int threads = 256;
  const int blocks = static_cast<int>(std::min<long long>((total + threads - 1) / threads, 148 * 8));
  pack_conv_kernel<<<blocks, threads, 0, st>>>(w, scale, wpk, wpkt, Cout, Cin, KK);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// FrozenBatchNorm2d fold (backbone.py:47-57): scale = w * rsqrt(rv + 1e-5); shift = b - rm * scale
__global__ void bn_fold_kernel(const float* w, const float* b, const float* rm, const float* rv, float* scale, float* shift, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float s = w[i] * rsqrtf(rv[i] + 1e-5f);
    scale[i] = s;
    shift[i] = b[i] - rm[i] * s;
  }
}
int bn_fold(const float* w, const float* b, const float* rm, const float* rv, float* scale, float* shift, int n, cudaStream_t st) {
  bn_fold_kernel<<<(n + 127) / 128, 128, 0, st>>>(w, b, rm, rv, scale, shift, n);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace act

namespace act {
__global__ void pack_conv1_kernel(const float* __restrict__ w, bf16* __restrict__ wpk, int KP) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * KP) return;
  const int co = i / KP, k = i % KP;
  wpk[i] = __float2bfloat16_rn(k < 147 ? w[co * 147 + k] : 0.f);
}
int pack_conv1_weights(const float* w, bf16* wpk, int KP, cudaStream_t st) {
  pack_conv1_kernel<<<(64 * KP + 255) / 256, 256, 0, st>>>(w, wpk, KP);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
}  // namespace act
