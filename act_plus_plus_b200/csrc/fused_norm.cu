// LayerNorm forward/backward (+positional add / bf16 casts), softmax forward/backward, column sums.
// All bandwidth-bound: one warp per row, 128-bit accesses, warp-shuffle reductions, no shared-memory round trips
// except the final cross-warp parameter-gradient reduction.
#include "kernels.h"

namespace act {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint2 pack4_bf16(float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 o;
  o.x = *reinterpret_cast<uint32_t*>(&lo);
  o.y = *reinterpret_cast<uint32_t*>(&hi);
  return o;
}

// ------------------------------------------------------------------------------------------ LayerNorm fwd
// y = LN(t) * gamma + beta (eps 1e-5, nn.LayerNorm as used at transformer.py:200-201,261-263).
// Optional extra outputs: bf16(y), bf16(y + pos[row % L]); mean / rstd for the backward.
template <int V>   // D = 128 * V
__global__ void ln_fwd_kernel(const float* __restrict__ t, const float* __restrict__ gamma, const float* __restrict__ beta,
                              const float* __restrict__ pos, int L, float* __restrict__ y, bf16* __restrict__ yb,
                              bf16* __restrict__ ypb, float* __restrict__ mean, float* __restrict__ rstd, int M) {
  pdl_trigger(); pdl_wait();
  constexpr int D = 128 * V;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= M) return;
  const float4* tp = reinterpret_cast<const float4*>(t + static_cast<long long>(row) * D);
  float4 x[V];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < V; ++j) { x[j] = tp[lane + 32 * j]; s += x[j].x + x[j].y + x[j].z + x[j].w; }
  const float mu = warp_sum(s) * (1.f / D);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const float a = x[j].x - mu, b = x[j].y - mu, c = x[j].z - mu, d = x[j].w - mu;
    q += a * a + b * b + c * c + d * d;
  }
  const float rs = rsqrtf(warp_sum(q) * (1.f / D) + 1e-5f);
  if (lane == 0) {
    if (mean) mean[row] = mu;
    if (rstd) rstd[row] = rs;
  }
  const float4* gp = reinterpret_cast<const float4*>(gamma);
  const float4* bp = reinterpret_cast<const float4*>(beta);
  const float4* pp = pos ? reinterpret_cast<const float4*>(pos + static_cast<long long>(row % L) * D) : nullptr;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = lane + 32 * j;
    const float4 g = __ldg(gp + c), b = __ldg(bp + c);
    float4 o;
    o.x = (x[j].x - mu) * rs * g.x + b.x; o.y = (x[j].y - mu) * rs * g.y + b.y;
    o.z = (x[j].z - mu) * rs * g.z + b.z; o.w = (x[j].w - mu) * rs * g.w + b.w;
    const long long off = static_cast<long long>(row) * D + 4 * c;
    if (y) *reinterpret_cast<float4*>(y + off) = o;
    if (yb) *reinterpret_cast<uint2*>(yb + off) = pack4_bf16(o.x, o.y, o.z, o.w);
    if (ypb) {
      const float4 p = pp ? __ldg(pp + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<uint2*>(ypb + off) = pack4_bf16(o.x + p.x, o.y + p.y, o.z + p.z, o.w + p.w);
    }
  }
}

int ln_fwd(const float* t, const float* gamma, const float* beta, const float* pos, int L, float* y, bf16* yb, bf16* ypb,
           float* mean, float* rstd, int M, int D, cudaStream_t st) {
  ACT_CHECK(D == 512 || D == 256, "LayerNorm width must be 256 or 512");
  if (M <= 0) return 0;
  const int wpb = 4;
  dim3 grid((M + wpb - 1) / wpb), block(32 * wpb);
  if (D == 512) ACT_CUDA(launch_pdl(ln_fwd_kernel<4>, dim3(grid), dim3(block), 0, st, t, gamma, beta, pos, L > 0 ? L : 1, y, yb, ypb, mean, rstd, M));
  else ACT_CUDA(launch_pdl(ln_fwd_kernel<2>, dim3(grid), dim3(block), 0, st, t, gamma, beta, pos, L > 0 ? L : 1, y, yb, ypb, mean, rstd, M));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ LayerNorm bwd
// dt = rstd * (dxh - mean(dxh) - xh * mean(dxh * xh)),  dxh = (dy + dy2) * gamma,  xh = (t - mean) * rstd
// dgamma += sum_rows dy * xh, dbeta += sum_rows dy  (fp32 atomics, one per column per block)
template <int V>
__global__ void ln_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ dy2, const float* __restrict__ t,
                              const float* __restrict__ mean, const float* __restrict__ rstd,
                              const float* __restrict__ gamma, float* __restrict__ dt, bf16* __restrict__ dtb,
                              float* __restrict__ dgamma, float* __restrict__ dbeta, int M, int rows_per_block,
                              const DropDesc drop) {
  pdl_trigger(); pdl_wait();
  constexpr int D = 128 * V;
  constexpr int WARPS = 4;
  __shared__ float red[2][WARPS][D];
  // train-mode dropout: t = x + dropout(h), so the bf16 copy (= the gradient handed to the producing linear layer) is
  // mask * dt / (1 - p) while the fp32 copy (the residual path) stays dt; the mask is regenerated, never stored
  const bool dropping = drop_on(drop);
  const uint32_t dstep = dropping ? __ldg(drop.step) : 0u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float4 ag[V], ab[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { ag[j] = make_float4(0.f, 0.f, 0.f, 0.f); ab[j] = ag[j]; }
  const float4* gp = reinterpret_cast<const float4*>(gamma);
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(M, r0 + rows_per_block);
  for (int row = r0 + warp; row < r1; row += WARPS) {
    const long long base = static_cast<long long>(row) * D;
    const float mu = mean[row], rs = rstd[row];
    float4 g[V], xh[V];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = lane + 32 * j;
      float4 d = *reinterpret_cast<const float4*>(dy + base + 4 * c);
      if (dy2) {
        const float4 e = *reinterpret_cast<const float4*>(dy2 + base + 4 * c);
        d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
      }
      const float4 tv = *reinterpret_cast<const float4*>(t + base + 4 * c);
      const float4 gm = __ldg(gp + c);
      xh[j] = make_float4((tv.x - mu) * rs, (tv.y - mu) * rs, (tv.z - mu) * rs, (tv.w - mu) * rs);
      ab[j].x += d.x; ab[j].y += d.y; ab[j].z += d.z; ab[j].w += d.w;
      ag[j].x += d.x * xh[j].x; ag[j].y += d.y * xh[j].y; ag[j].z += d.z * xh[j].z; ag[j].w += d.w * xh[j].w;
      g[j] = make_float4(d.x * gm.x, d.y * gm.y, d.z * gm.z, d.w * gm.w);
      s1 += g[j].x + g[j].y + g[j].z + g[j].w;
      s2 += g[j].x * xh[j].x + g[j].y * xh[j].y + g[j].z * xh[j].z + g[j].w * xh[j].w;
    }
    s1 = warp_sum(s1) * (1.f / D);
    s2 = warp_sum(s2) * (1.f / D);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = lane + 32 * j;
      float4 o;
      o.x = rs * (g[j].x - s1 - xh[j].x * s2); o.y = rs * (g[j].y - s1 - xh[j].y * s2);
      o.z = rs * (g[j].z - s1 - xh[j].z * s2); o.w = rs * (g[j].w - s1 - xh[j].w * s2);
      if (dt) *reinterpret_cast<float4*>(dt + base + 4 * c) = o;
      if (dtb) {
        if (dropping) {
          const unsigned long long e = static_cast<unsigned long long>(base) + 4 * c;
          const uint32_t m = drop_keep8(drop, e >> 3, dstep) >> (e & 7);
          o.x = (m & 1u) ? o.x * drop.scale : 0.f; o.y = (m & 2u) ? o.y * drop.scale : 0.f;
          o.z = (m & 4u) ? o.z * drop.scale : 0.f; o.w = (m & 8u) ? o.w * drop.scale : 0.f;
        }
        *reinterpret_cast<uint2*>(dtb + base + 4 * c) = pack4_bf16(o.x, o.y, o.z, o.w);
      }
    }
  }
  if (dgamma == nullptr) return;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = 4 * (lane + 32 * j);
    red[0][warp][c] = ag[j].x; red[0][warp][c + 1] = ag[j].y; red[0][warp][c + 2] = ag[j].z; red[0][warp][c + 3] = ag[j].w;
    red[1][warp][c] = ab[j].x; red[1][warp][c + 1] = ab[j].y; red[1][warp][c + 2] = ab[j].z; red[1][warp][c + 3] = ab[j].w;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) { a += red[0][w][c]; b += red[1][w][c]; }
    atomicAdd(dgamma + c, a);
    atomicAdd(dbeta + c, b);
  }
}

int ln_bwd(const float* dy, const float* dy2, const float* t, const float* mean, const float* rstd, const float* gamma,
           float* dt, bf16* dtb, float* dgamma, float* dbeta, int M, int D, cudaStream_t st, const DropDesc* dropp) {
  const DropDesc drop = dropp ? *dropp : DropDesc();
  ACT_CHECK(D == 512 || D == 256, "LayerNorm width must be 256 or 512");
  if (M <= 0) return 0;
  int rpb = 16;
  while (rpb > 4 && (M + rpb - 1) / rpb < 148 * 2) rpb >>= 1;
  dim3 grid((M + rpb - 1) / rpb), block(128);
  if (D == 512) ACT_CUDA(launch_pdl(ln_bwd_kernel<4>, dim3(grid), dim3(block), 0, st, dy, dy2, t, mean, rstd, gamma, dt, dtb, dgamma, dbeta, M, rpb, drop));
  else ACT_CUDA(launch_pdl(ln_bwd_kernel<2>, dim3(grid), dim3(block), 0, st, dy, dy2, t, mean, rstd, gamma, dt, dtb, dgamma, dbeta, M, rpb, drop));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ x (+pos) -> bf16
// xb = bf16(x), xpb = bf16(x + pos[row % L])   (with_pos_embed of transformer.py:205-206 feeding q/k only)
__global__ void add_pos_cast_kernel(const float* __restrict__ x, const float* __restrict__ pos, int L, int D4,
                                    bf16* __restrict__ xb, bf16* __restrict__ xpb, long long n4) {
  pdl_trigger(); pdl_wait();
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * 1LL * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    if (xb) reinterpret_cast<uint2*>(xb)[i] = pack4_bf16(v.x, v.y, v.z, v.w);
    if (xpb) {
      float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
      if (pos) {
        const long long row = i / D4;
        p = __ldg(reinterpret_cast<const float4*>(pos) + (row % L) * D4 + (i % D4));
      }
      reinterpret_cast<uint2*>(xpb)[i] = pack4_bf16(v.x + p.x, v.y + p.y, v.z + p.z, v.w + p.w);
    }
  }
}
int add_pos_cast(const float* x, const float* pos, int L, bf16* xb, bf16* xpb, int M, int D, cudaStream_t st) {
  ACT_CHECK(D % 4 == 0, "D must be a multiple of 4");
  const long long n4 = 1LL * M * D / 4;
  if (n4 <= 0) return 0;
  const int blocks = static_cast<int>(std::min<long long>((n4 + 255) / 256, 148 * 8));
  ACT_CUDA(launch_pdl(add_pos_cast_kernel, dim3(blocks), dim3(256), 0, st, x, pos, L > 0 ? L : 1, D / 4, xb, xpb, n4));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ softmax fwd
// P[r, :L] = softmax(S[r, :L] masked by key_padding_mask[b, :]) in bf16, P[r, L:Lp] = 0.
// One warp per row; rows R = B * heads * Lq, b = r / rows_per_batch.  Lp <= 1024.
__global__ void softmax_fwd_kernel(const float* __restrict__ S, const uint8_t* __restrict__ mask, bf16* __restrict__ P,
                                   int R, int L, int Lp, int rows_per_batch) {
  pdl_trigger(); pdl_wait();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= R) return;
  const float* s = S + static_cast<long long>(row) * Lp;
  const uint8_t* m = mask ? mask + static_cast<long long>(row / rows_per_batch) * L : nullptr;
  constexpr int MAXV = 32;                 // 32 lanes * 32 = 1024 columns
  float v[MAXV];
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + 32 * j;
    float x = -INFINITY;
    if (c < L && !(m && m[c])) x = s[c];
    v[j] = x;
    mx = fmaxf(mx, x);
  }
  mx = warp_max(mx);
  if (mx == -INFINITY) mx = 0.f;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) { v[j] = __expf(v[j] - mx); sum += v[j]; }
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  bf16* p = P + static_cast<long long>(row) * Lp;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + 32 * j;
    if (c < Lp) p[c] = __float2bfloat16_rn(c < L ? v[j] * inv : 0.f);
  }
}
int softmax_fwd(const float* S, const uint8_t* mask, bf16* P, int R, int L, int Lp, int rows_per_batch, cudaStream_t st) {
  ACT_CHECK(Lp <= 1024 && L <= Lp, "softmax row too long (max 1024 keys)");
  if (R <= 0) return 0;
  const int wpb = 4;
  ACT_CUDA(launch_pdl(softmax_fwd_kernel, dim3((R + wpb - 1) / wpb), dim3(32 * wpb), 0, st, S, mask, P, R, L, Lp, rows_per_batch > 0 ? rows_per_batch : 1));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// dS = scale * P * (dP - sum_j P*dP)  -> bf16, pad columns zero
__global__ void softmax_bwd_kernel(const bf16* __restrict__ P, const float* __restrict__ dP, bf16* __restrict__ dS,
                                   int R, int L, int Lp, float scale) {
  pdl_trigger(); pdl_wait();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= R) return;
  const bf16* p = P + static_cast<long long>(row) * Lp;
  const float* d = dP + static_cast<long long>(row) * Lp;
  constexpr int MAXV = 32;
  float pv[MAXV], dv[MAXV];
  float dot = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + 32 * j;
    pv[j] = 0.f; dv[j] = 0.f;
    if (c < L) { pv[j] = __bfloat162float(p[c]); dv[j] = d[c]; }
    dot += pv[j] * dv[j];
  }
  dot = warp_sum(dot);
  bf16* o = dS + static_cast<long long>(row) * Lp;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + 32 * j;
    if (c < Lp) o[c] = __float2bfloat16_rn(c < L ? scale * pv[j] * (dv[j] - dot) : 0.f);
  }
}
int softmax_bwd(const bf16* P, const float* dP, bf16* dS, int R, int L, int Lp, float scale, cudaStream_t st) {
  ACT_CHECK(Lp <= 1024 && L <= Lp, "softmax row too long (max 1024 keys)");
  if (R <= 0) return 0;
  const int wpb = 4;
  ACT_CUDA(launch_pdl(softmax_bwd_kernel, dim3((R + wpb - 1) / wpb), dim3(32 * wpb), 0, st, P, dP, dS, R, L, Lp, scale));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ column sums
// out[n] += sum_m X[m*ld + n]  (bias gradients).  Block = 32 x 8 threads: 64 columns (bf16x2 per thread) x row stripes.
__global__ void colsum_bf16_kernel(const bf16* __restrict__ X, long long ld, int M, int N, float* __restrict__ out,
                                   int rows_per_block) {
  pdl_trigger(); pdl_wait();
  __shared__ float2 red[8][32];
  const int c = blockIdx.x * 64 + threadIdx.x * 2;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float2 acc = make_float2(0.f, 0.f);
  if (c < N) {
    for (int r = r0 + threadIdx.y; r < r1; r += 8) {
      const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(X + r * ld + c);
      const float2 f = __bfloat1622float2(v);
      acc.x += f.x; acc.y += f.y;
    }
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < N) {
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) { s.x += red[i][threadIdx.x].x; s.y += red[i][threadIdx.x].y; }
    atomicAdd(out + c, s.x);
    if (c + 1 < N) atomicAdd(out + c + 1, s.y);
  }
}
int colsum_bf16(const bf16* X, long long ld, int M, int N, float* out, cudaStream_t st) {
  ACT_CHECK(N % 2 == 0 && ld % 2 == 0, "colsum needs even N / ld");
  if (M <= 0 || N <= 0) return 0;
  const int cb = (N + 63) / 64;
  int rpb = 256;
  while (rpb > 32 && 1LL * cb * ((M + rpb - 1) / rpb) < 148 * 2) rpb >>= 1;
  dim3 grid(cb, (M + rpb - 1) / rpb), block(32, 8);
  ACT_CUDA(launch_pdl(colsum_bf16_kernel, dim3(grid), dim3(block), 0, st, X, ld, M, N, out, rpb));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace act
