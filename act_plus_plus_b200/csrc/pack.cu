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

// Batched re-pack of every conv of a backbone in ONE launch (runs after each AdamW step).  A block owns a 32(co) x
// 32(ci) x KK tile: coalesced fp32 reads of w[co][ci0..ci0+32][*] into shared memory, then 64-byte bf16 runs along ci
// (wpk) and along co (wpkt).  KK is a template constant: with a runtime divisor in the index arithmetic this kernel was
// instruction-bound (110 us for 11 M weights).
template <int KK>
__device__ __forceinline__ void pack_conv_body(const PackJob& jb, int b, float (*tile)[32 * 9 + 1]) {
  const int ci_tiles = (jb.Cin + 31) / 32;
  const int co0 = (b / ci_tiles) * 32, ci0 = (b % ci_tiles) * 32;
  constexpr int run = 32 * KK;                               // floats per co row of the tile (contiguous in w)
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int valid = min(32, jb.Cin - ci0) * KK;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int co = wrp + 8 * c;
    const bool ok = co0 + co < jb.Cout;
    const float* wrow = jb.w + (1LL * (co0 + co) * jb.Cin + ci0) * KK;
#pragma unroll
    for (int o = lane; o < run; o += 32) tile[co][o] = (ok && o < valid) ? __ldg(wrow + o) : 0.f;
  }
  __syncthreads();
  const bool ci_ok = ci0 + lane < jb.Cin;
#pragma unroll
  for (int c = 0; c < 4; ++c) {                              // wpk[co][t*Cin + ci]: ci = lane fastest
    const int co = wrp + 8 * c;
    if (co0 + co >= jb.Cout || !ci_ok) continue;
    bf16* orow = jb.wpk + (1LL * (co0 + co) * KK) * jb.Cin + ci0 + lane;
#pragma unroll
    for (int t = 0; t < KK; ++t) orow[1LL * t * jb.Cin] = __float2bfloat16_rn(tile[co][lane * KK + t]);
  }
  if (jb.wpkt) {                                             // wpkt[ci][t*Cout + co]: co = lane fastest
    const bool co_ok = co0 + lane < jb.Cout;
    const float sc = (jb.scale && co_ok) ? jb.scale[co0 + lane] : 1.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int ci = wrp + 8 * c;
      if (ci0 + ci >= jb.Cin || !co_ok) continue;
      bf16* orow = jb.wpkt + (1LL * (ci0 + ci) * KK) * jb.Cout + co0 + lane;
#pragma unroll
      for (int t = 0; t < KK; ++t) orow[1LL * t * jb.Cout] = __float2bfloat16_rn(tile[lane][ci * KK + t] * sc);
    }
  }
}
__global__ void __launch_bounds__(256) pack_conv_batched_kernel(const __grid_constant__ PackJobs jobs) {
  pdl_trigger(); pdl_wait();
  __shared__ float tile[32][32 * 9 + 1];
  int b = blockIdx.x, j = 0;
  while (j + 1 < jobs.n && b >= jobs.job[j + 1].tile0) ++j;
  const PackJob& jb = jobs.job[j];
  b -= jb.tile0;
  if (jb.KK == 9) pack_conv_body<9>(jb, b, tile);
  else pack_conv_body<1>(jb, b, tile);
}
int pack_conv_batched(PackJobs& jobs, cudaStream_t st) {
  ACT_CHECK(jobs.n >= 1 && jobs.n <= PackJobs::MAX, "pack_conv_batched: 1..64 jobs");
  int tiles = 0;
  for (int i = 0; i < jobs.n; ++i) {
    ACT_CHECK(jobs.job[i].KK == 1 || jobs.job[i].KK == 9, "pack_conv_batched: 1x1 or 3x3 kernels");
    jobs.job[i].tile0 = tiles;
    tiles += ((jobs.job[i].Cout + 31) / 32) * ((jobs.job[i].Cin + 31) / 32);
  }
  ACT_CUDA(launch_pdl(pack_conv_batched_kernel, dim3(tiles), dim3(256), 0, st, jobs));
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

namespace act {
// Stem (7x7/s2, 3 -> 64) weights for the implicit-GEMM formulation: one k-block per kernel row kh, 64 columns =
// 16 pixels x 4 channels of the padded NHWC4 image window; column kw*4 + c holds w[co][c][kh][kw] (kw < 7, c < 3).
// Rows 64*j .. 64*j+63 (j = 1, 2, 3) hold the same filters shifted right by j output pixels (2j input pixels = 8j
// columns; the last real column is 24 + 27 = 51 < 64): with them one GEMM row produces up to FOUR horizontally adjacent
// output pixels (N = 256 instead of 64 at the same MMA count per row), reading the window that starts at the first one.
__global__ void pack_stem_kernel(const float* __restrict__ w, bf16* __restrict__ wpk) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 256 * 448) return;
  const int row = i / 448, co = row & 63, kh = (i % 448) / 64, e = i % 64, kw = (e >> 2) - 2 * (row >> 6), c = e & 3;
  wpk[i] = __float2bfloat16_rn((kw >= 0 && kw < 7 && c < 3) ? w[((co * 3 + c) * 7 + kh) * 7 + kw] : 0.f);
}
int pack_stem_weights(const float* w, bf16* wpk, cudaStream_t st) {
  pack_stem_kernel<<<(256 * 448 + 255) / 256, 256, 0, st>>>(w, wpk);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// Paired 3x3/s1 weights for 64-channel convs (conv_ops.cu::pair_launch): the conv over pixel PAIRS [N,H,W/2,128] has
// Cin' = Cout' = 128; dst [128][9 taps * 128]; block sp = tap*2 + pi is the 64 input channels of pixel parity pi of the
// pair at (dh, dwp).  Rows 0..63 (even output pixel) take source slot even_slot[sp] of the packed [64][9*64] operand,
// rows 64..127 (odd output pixel) odd_slot[sp]; -1 = zero block.
__global__ void pack_pair_kernel(PairJobs jobs) {
  const PairJob& jb = jobs.job[blockIdx.y];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 128 * 18 * 8) return;
  const int e8 = i % 8, sp = (i / 8) % 18, row = i / 144;
  const int slot = row < 64 ? jb.even_slot[sp] : jb.odd_slot[sp];
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (slot >= 0) v = *reinterpret_cast<const uint4*>(jb.src + (1LL * (row & 63) * 9 + slot) * 64 + e8 * 8);
  *reinterpret_cast<uint4*>(jb.dst + (1LL * row * 18 + sp) * 64 + e8 * 8) = v;
}
int pack_pair_batched(const PairJobs& jobs, cudaStream_t st) {
  if (jobs.n <= 0) return 0;
  pack_pair_kernel<<<dim3((128 * 18 * 8 + 255) / 256, jobs.n), 256, 0, st>>>(jobs);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
}  // namespace act
