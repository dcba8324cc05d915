// Small fused kernels of the ACT path: strided fp32 mini-GEMM (the K<=32 / N<=64 projections and their gradients),
// L1+KL loss with its gradients, reparametrisation, max-pool fwd/bwd.
#include "kernels.h"

namespace act {

// ------------------------------------------------------------------------------------------ mini GEMM (fp32)
// C(i,j) (+)= alpha * sum_r A(i,r) * B(r,j) (+ bias[j]); every index uses a batched-affine address
// off(idx) = (idx / rpb) * bs + (idx % rpb) * s, so token-buffer rows, transposes and broadcasts need no copies.
__device__ __forceinline__ long long ioff(const IdxMap& m, int idx) {
  return static_cast<long long>(idx / m.rpb) * m.bs + static_cast<long long>(idx % m.rpb) * m.s;
}
// Block = 32 output columns x 8 reduction slices: thread (tx, ty) accumulates r = r0+ty, r0+ty+8, ... for column j,
// the 8 partial sums meet in shared memory.  grid.z splits long reductions further (atomic accumulate).
__global__ void __launch_bounds__(256) sgemm_generic_kernel(SgemmArgs a) {
  pdl_trigger(); pdl_wait();
  __shared__ float part[8][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int j = blockIdx.x * 32 + tx;
  const int i = blockIdx.y;
  const int rchunk = (a.R + gridDim.z - 1) / gridDim.z;
  const int r0 = blockIdx.z * rchunk, r1 = min(a.R, r0 + rchunk);
  float acc = 0.f;
  if (j < a.J) {
    const float* Ai = a.A + ioff(a.ai, i);
    const float* Bj = a.B + ioff(a.bj, j);
    const bool flat = a.ar.rpb >= a.R && a.br.rpb >= a.R;       // no batched index maps on the reduction axis
    if (flat) {
      const long long sa = a.ar.s, sb = a.br.s;
      int r = r0 + ty;
      for (; r + 24 < r1; r += 32) {
        const float a0 = Ai[r * sa], a1 = Ai[(r + 8) * sa], a2 = Ai[(r + 16) * sa], a3 = Ai[(r + 24) * sa];
        const float b0 = Bj[r * sb], b1 = Bj[(r + 8) * sb], b2 = Bj[(r + 16) * sb], b3 = Bj[(r + 24) * sb];
        acc += a0 * b0 + a1 * b1 + a2 * b2 + a3 * b3;
      }
      for (; r < r1; r += 8) acc += Ai[r * sa] * Bj[r * sb];
    } else {
      int r = r0 + ty;
      for (; r + 24 < r1; r += 32) {                             // four independent load pairs in flight
        const float a0 = Ai[ioff(a.ar, r)], a1 = Ai[ioff(a.ar, r + 8)], a2 = Ai[ioff(a.ar, r + 16)], a3 = Ai[ioff(a.ar, r + 24)];
        const float b0 = Bj[ioff(a.br, r)], b1 = Bj[ioff(a.br, r + 8)], b2 = Bj[ioff(a.br, r + 16)], b3 = Bj[ioff(a.br, r + 24)];
        acc += a0 * b0 + a1 * b1 + a2 * b2 + a3 * b3;
      }
      for (; r < r1; r += 8) acc += Ai[ioff(a.ar, r)] * Bj[ioff(a.br, r)];
    }
  }
  part[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && j < a.J) {
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) v += part[q][tx];
    v *= a.alpha;
    if (a.bias && blockIdx.z == 0) v += a.bias[j];
    float* c = a.C + ioff(a.ci, i) + ioff(a.cj, j);
    if (a.accumulate || gridDim.z > 1) atomicAdd(c, v);
    else *c = v;
  }
}
// Short reductions (R <= 64, e.g. the 14/16-wide input projections and d_hs = d_a W): one thread per output, lanes
// along j, the A row is a warp broadcast.  The block shape of the generic kernel would leave 7/8 of its threads idle.
__global__ void __launch_bounds__(256) sgemm_short_kernel(SgemmArgs a) {
  pdl_trigger(); pdl_wait();
  const int j = blockIdx.x * 256 + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= a.J) return;
  const float* Ai = a.A + ioff(a.ai, i);
  const float* Bj = a.B + ioff(a.bj, j);
  const long long sa = a.ar.s, sb = a.br.s;
  float acc = 0.f;
  if (a.a_mean) {
    for (int r = 0; r < a.R; ++r) acc += ((__ldg(Ai + r * sa) - __ldg(a.a_mean + r)) / __ldg(a.a_std + r)) * __ldg(Bj + r * sb);
  } else {
#pragma unroll 4
    for (int r = 0; r < a.R; ++r) acc += __ldg(Ai + r * sa) * __ldg(Bj + r * sb);
  }
  acc *= a.alpha;
  if (a.bias) acc += a.bias[j];
  float* c = a.C + ioff(a.ci, i) + ioff(a.cj, j);
  if (a.accumulate) atomicAdd(c, acc);
  else *c = acc;
}
// Long reductions with both operands contiguous along r and few output columns (action / is_pad heads, latent_proj):
// one warp per (output row, group of 4 columns), lanes along r (coalesced); 4 independent dot products keep enough
// loads in flight to cover the L2 latency.
__global__ void __launch_bounds__(256) sgemm_rowdot_kernel(SgemmArgs a) {
  pdl_trigger(); pdl_wait();
  const int lane = threadIdx.x & 31;
  const int jg = (a.J + 3) / 4;
  const int w = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= a.I * jg) return;
  const int i = w / jg, j0 = (w % jg) * 4;
  const float* Ai = a.A + ioff(a.ai, i);
  const float* Bj[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) Bj[q] = a.B + ioff(a.bj, min(j0 + q, a.J - 1));
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int r = lane; r < a.R; r += 32) {
    const float av = __ldg(Ai + r);
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[q] += av * __ldg(Bj[q] + r);
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
  }
  if (lane < 4 && j0 + lane < a.J) {
    const int j = j0 + lane;
    float v = (lane == 0 ? acc[0] : lane == 1 ? acc[1] : lane == 2 ? acc[2] : acc[3]) * a.alpha;
    if (a.bias) v += a.bias[j];
    float* c = a.C + ioff(a.ci, i) + ioff(a.cj, j);
    if (a.accumulate) atomicAdd(c, v);
    else *c = v;
  }
}
int sgemm_generic(const SgemmArgs& a, cudaStream_t st) {
  if (a.I <= 0 || a.J <= 0) return 0;
  ACT_CHECK(a.I <= 65535, "sgemm_generic: too many rows");
  const bool flat = a.ar.rpb >= a.R && a.br.rpb >= a.R;
  ACT_CHECK(!a.a_mean || (a.a_std && flat && a.R <= 64 && a.J >= 128), "sgemm: operand z-score is implemented for the short-reduction form only");
  if (flat && a.R <= 64 && a.J >= 128) {
    ACT_CUDA(launch_pdl(sgemm_short_kernel, dim3((a.J + 255) / 256, a.I), dim3(256), 0, st, a));
    ACT_CUDA(cudaGetLastError());
    return 0;
  }
  if (flat && a.ar.s == 1 && a.br.s == 1 && a.J <= 64 && a.R >= 128) {
    ACT_CUDA(launch_pdl(sgemm_rowdot_kernel, dim3((a.I * ((a.J + 3) / 4) + 7) / 8), dim3(256), 0, st, a));
    ACT_CUDA(cudaGetLastError());
    return 0;
  }
  const int jb = (a.J + 31) / 32;
  int split = 1;
  if (a.accumulate && a.R > 64) {
    // enough blocks to fill the chip, at least 32 reduction steps per block
    const long long blocks = 1LL * jb * a.I;
    split = static_cast<int>(std::max<long long>(1, std::min<long long>((a.R + 31) / 32, (2 * 148 + blocks - 1) / blocks)));
    // these launches sit at the very end of a dependency chain (token-projection gradients): a block is latency-bound, so
    // no thread should walk more than ~8 dependent-latency steps even when the chip is already "full"
    split = static_cast<int>(std::max<long long>(split, std::min<long long>((a.R + 63) / 64, 32)));
  }
  dim3 grid(jb, a.I, split);
  ACT_CUDA(launch_pdl(sgemm_generic_kernel, grid, dim3(32, 8), 0, st, a));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ loss
// policy.py:314-318 + 378-391:  l1 = mean over B*nq*A of |a - a_hat| * ~pad ; kl = mean_b sum_z -0.5(1+lv-mu^2-e^lv)
// loss = l1 + kl_weight * kl.  Also emits dL/da_hat, and the KL part of dL/dmu, dL/dlogvar (scaled by `gscale`).
__global__ void act_loss_kernel(const float* __restrict__ a_hat, const float* __restrict__ actions, long long act_bs,
                                const uint8_t* __restrict__ is_pad, long long pad_bs, const float* __restrict__ mu,
                                const float* __restrict__ logvar, int B, int nq, int A, int Z, float kl_weight,
                                float gscale, float* __restrict__ losses, float* __restrict__ d_a_hat,
                                float* __restrict__ d_mu, float* __restrict__ d_logvar, const float* __restrict__ gvec) {
  // gvec (device, optional) = the incoming gradient of the three returned scalars {l1, kl, loss}: any of them may be the
  // root of loss.backward() (policy.py:316-318 returns all three); loss = l1 + kl_weight * kl
  const float gl = gvec ? gvec[0] + gvec[2] : gscale;
  const float gk = gvec ? gvec[1] + kl_weight * gvec[2] : gscale * kl_weight;
  __shared__ float red[32];
  __shared__ float tot[2];
  const int tid = threadIdx.x, nthr = blockDim.x;
  const long long n = 1LL * B * nq * A;
  const float inv_n = 1.f / static_cast<float>(n);
  float s = 0.f;
  for (long long i = tid; i < n; i += nthr) {
    const int a = static_cast<int>(i % A);
    const long long r = i / A;
    const int q = static_cast<int>(r % nq), b = static_cast<int>(r / nq);
    const float diff = a_hat[i] - actions[b * act_bs + static_cast<long long>(q) * A + a];
    const bool pad = is_pad[b * pad_bs + q] != 0;
    if (!pad) s += fabsf(diff);
    if (d_a_hat) d_a_hat[i] = pad ? 0.f : gl * inv_n * ((diff > 0.f) ? 1.f : (diff < 0.f ? -1.f : 0.f));
  }
  float k = 0.f;
  for (int i = tid; i < B * Z; i += nthr) {
    const float m = mu[i], lv = logvar[i], e = __expf(lv);
    k += -0.5f * (1.f + lv - m * m - e);
    if (d_mu) d_mu[i] = gk * m / B;
    if (d_logvar) d_logvar[i] = gk * 0.5f * (e - 1.f) / B;
  }
  for (int pass = 0; pass < 2; ++pass) {
    float v = pass == 0 ? s : k;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    if (tid < 32) {
      float w = tid < (nthr >> 5) ? red[tid] : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
      if (tid == 0) tot[pass] = w;
    }
    __syncthreads();
  }
  if (tid == 0) {
    const float l1 = tot[0] * inv_n, kl = tot[1] / B;
    losses[0] = l1; losses[1] = kl; losses[2] = l1 + kl_weight * kl;
  }
}
int act_loss(const float* a_hat, const float* actions, long long act_bs, const uint8_t* is_pad, long long pad_bs,
             const float* mu, const float* logvar, int B, int nq, int A, int Z, float kl_weight, float gscale,
             float* losses, float* d_a_hat, float* d_mu, float* d_logvar, cudaStream_t st, const float* gvec) {
  act_loss_kernel<<<1, 1024, 0, st>>>(a_hat, actions, act_bs, is_pad, pad_bs, mu, logvar, B, nq, A, Z, kl_weight, gscale,
                                      losses, d_a_hat, d_mu, d_logvar, gvec);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ latent
// detr_vae.py:19-22,147-150: mu = info[:, :Z], logvar = info[:, Z:], z = mu + exp(logvar/2) * eps
__global__ void reparam_fwd_kernel(const float* info, const float* eps, float* mu, float* logvar, float* z, int B, int Z) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * Z) return;
  const int b = i / Z, k = i % Z;
  const float m = info[b * 2 * Z + k], lv = info[b * 2 * Z + Z + k];
  mu[i] = m; logvar[i] = lv;
  z[i] = m + __expf(0.5f * lv) * eps[i];
}
int reparam_fwd(const float* info, const float* eps, float* mu, float* logvar, float* z, int B, int Z, cudaStream_t st) {
  reparam_fwd_kernel<<<(B * Z + 127) / 128, 128, 0, st>>>(info, eps, mu, logvar, z, B, Z);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
// d_info[:, :Z] = d_mu_kl + dz ; d_info[:, Z:] = d_lv_kl + dz * eps * 0.5 * exp(lv/2)
__global__ void reparam_bwd_kernel(const float* dz, const float* eps, const float* logvar, const float* d_mu_kl,
                                   const float* d_lv_kl, float* d_info, int B, int Z) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * Z) return;
  const int b = i / Z, k = i % Z;
  d_info[b * 2 * Z + k] = d_mu_kl[i] + dz[i];
  d_info[b * 2 * Z + Z + k] = d_lv_kl[i] + dz[i] * eps[i] * 0.5f * __expf(0.5f * logvar[i]);
}
int reparam_bwd(const float* dz, const float* eps, const float* logvar, const float* d_mu_kl, const float* d_lv_kl,
                float* d_info, int B, int Z, cudaStream_t st) {
  reparam_bwd_kernel<<<(B * Z + 127) / 128, 128, 0, st>>>(dz, eps, logvar, d_mu_kl, d_lv_kl, d_info, B, Z);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ max-pool 3x3 / s2 / p1
// NHWC bf16, 8 channels (16 B) per thread.  Forward also records the arg-max tap (first max in kh,kw scan order,
// as torch) so the backward is a gather.
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, int N,
                                                          int H, int W, int C, int Ho, int Wo) {
  // one thread = 8 channels of TWO horizontally adjacent outputs: their windows share one input column, so a pair costs
  // 15 loads instead of 18, and all of them are issued before the first compare
  pdl_trigger(); pdl_wait();
  const int C8 = C / 8, Wo2 = (Wo + 1) / 2;
  const long long total = 1LL * N * Ho * Wo2 * C8;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long r = i / C8;
    const int ox = 2 * static_cast<int>(r % Wo2); r /= Wo2;
    const int oy = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    uint4 v[3][5];
    bool ok[3][5];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int iy = 2 * oy + kh - 1;
#pragma unroll
      for (int kc = 0; kc < 5; ++kc) {                       // columns 2*ox-1 .. 2*ox+3; column 2*ox+1 feeds both outputs
        const int ix = 2 * ox + kc - 1;
        ok[kh][kc] = iy >= 0 && iy < H && ix >= 0 && ix < W;
        if (ok[kh][kc]) v[kh][kc] = *reinterpret_cast<const uint4*>(x + ((1LL * n * H + iy) * W + ix) * C + c8 * 8);
      }
    }
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      if (ox + o >= Wo) break;
      float best[8];
      uint8_t bi[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) { best[e] = -INFINITY; bi[e] = 0; }
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {                     // scan order kh, kw: the first maximum wins, as torch
          const int kc = 2 * o + kw;
          if (!ok[kh][kc]) continue;
          const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v[kh][kc]);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = __bfloat1622float2(h2[e]);
            if (f.x > best[2 * e]) { best[2 * e] = f.x; bi[2 * e] = kh * 3 + kw; }
            if (f.y > best[2 * e + 1]) { best[2 * e + 1] = f.y; bi[2 * e + 1] = kh * 3 + kw; }
          }
        }
      uint4 ov;
      __nv_bfloat162* o2 = reinterpret_cast<__nv_bfloat162*>(&ov);
#pragma unroll
      for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(best[2 * e], best[2 * e + 1]);
      const long long ooff = ((1LL * n * Ho + oy) * Wo + ox + o) * C + c8 * 8;
      *reinterpret_cast<uint4*>(y + ooff) = ov;
      // tap 9 = "no tap": the window maximum is not positive, so the ReLU fused into the backward passes nothing; the
      // backward then needs no look at x at all
#pragma unroll
      for (int e = 0; e < 8; ++e) if (!(best[e] > 0.f)) bi[e] = 9;
      uint2 p;
      p.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
      p.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
      *reinterpret_cast<uint2*>(idx + ooff) = p;
    }
  }
}
// Inference form (no arg-max record): packed bf16x2 maxima, one thread = 8 channels of TWO horizontally adjacent
// outputs (their windows share one column: 15 loads for 2 outputs instead of 18).
__global__ void maxpool_fwd_noidx_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int N, int H, int W, int C,
                                         int Ho, int Wo) {
  pdl_trigger(); pdl_wait();
  const int C8 = C / 8, Wo2 = (Wo + 1) / 2;
  const long long total = 1LL * N * Ho * Wo2 * C8;
  const __nv_bfloat162 ninf = __halves2bfloat162(__ushort_as_bfloat16(0xFF80), __ushort_as_bfloat16(0xFF80));
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long r = i / C8;
    const int ox = 2 * static_cast<int>(r % Wo2); r /= Wo2;
    const int oy = static_cast<int>(r % Ho);
    const int n = static_cast<int>(r / Ho);
    __nv_bfloat162 a[4], b[4];                               // outputs ox and ox + 1
#pragma unroll
    for (int e = 0; e < 4; ++e) { a[e] = ninf; b[e] = ninf; }
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int iy = 2 * oy + kh - 1;
      if (iy < 0 || iy >= H) continue;
      const bf16* row = x + ((1LL * n * H + iy) * W) * C + c8 * 8;
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) {                       // columns 2*ox-1 .. 2*ox+3; column 2*ox+1 feeds both outputs
        const int ix = 2 * ox + kw - 1;
        if (ix < 0 || ix >= W) continue;
        const uint4 v = *reinterpret_cast<const uint4*>(row + 1LL * ix * C);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (kw <= 2) a[e] = __hmax2(a[e], h2[e]);
          if (kw >= 2) b[e] = __hmax2(b[e], h2[e]);
        }
      }
    }
    const long long ooff = ((1LL * n * Ho + oy) * Wo + ox) * C + c8 * 8;
    *reinterpret_cast<uint4*>(y + ooff) = *reinterpret_cast<const uint4*>(a);
    if (ox + 1 < Wo) *reinterpret_cast<uint4*>(y + ooff + C) = *reinterpret_cast<const uint4*>(b);
  }
}

int maxpool_fwd(const bf16* x, bf16* y, uint8_t* idx, int N, int H, int W, int C, cudaStream_t st) {
  ACT_CHECK(C % 8 == 0, "maxpool needs C % 8 == 0");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = 1LL * N * Ho * ((Wo + 1) / 2) * (C / 8);       // both kernels: two adjacent outputs per thread
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  if (idx) ACT_CUDA(launch_pdl(maxpool_fwd_kernel, dim3(blocks), dim3(256), 0, st, x, y, idx, N, H, W, C, Ho, Wo));
  else ACT_CUDA(launch_pdl(maxpool_fwd_noidx_kernel, dim3(blocks), dim3(256), 0, st, x, y, N, H, W, C, Ho, Wo));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// dx[n,iy,ix,c] = (x > 0 ? 1 : 0) * sum over windows whose arg-max is (iy,ix) of dy   (ReLU of the conv1 block fused).
// One thread = 8 channels of a 2x2 block of input pixels (rows 2k, 2k+1; columns 2j, 2j+1): exactly the four windows
// (k..k+1, j..j+1) reach it, so their arg-max records and gradients are loaded once, all eight loads in flight together
// (the per-pixel form walked up to four windows one dependent round trip at a time and was latency-bound: 71 us for a
// 190 MB pass).  x may be NULL: the forward records tap 9 for windows whose maximum is not positive, which is the only
// case the ReLU mask changes, so the 79 MB read of x is skipped on the engine's path.
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          const bf16* __restrict__ x, bf16* __restrict__ dx, int N, int H,
                                                          int W, int C, int Ho, int Wo) {
  pdl_trigger(); pdl_wait();
  const int C8 = C / 8, H2 = (H + 1) / 2, W2 = (W + 1) / 2;
  const long long total = 1LL * N * H2 * W2 * C8;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    const int c8 = static_cast<int>(i % C8);
    long long r = i / C8;
    const int j = static_cast<int>(r % W2); r /= W2;
    const int k = static_cast<int>(r % H2);
    const int n = static_cast<int>(r / H2);
    uint2 rec[2][2];
    uint4 g[2][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const bool ok = k + a < Ho && j + b < Wo;
        const long long ooff = ((1LL * n * Ho + min(k + a, Ho - 1)) * Wo + min(j + b, Wo - 1)) * C + c8 * 8;
        rec[a][b] = ok ? *reinterpret_cast<const uint2*>(idx + ooff) : make_uint2(0x0f0f0f0fu, 0x0f0f0f0fu);
        g[a][b] = ok ? *reinterpret_cast<const uint4*>(dy + ooff) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
    for (int py = 0; py < 2; ++py) {
      const int iy = 2 * k + py;
      if (iy >= H) continue;
#pragma unroll
      for (int px = 0; px < 2; ++px) {
        const int ix = 2 * j + px;
        if (ix >= W) continue;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        // window (k+a, j+b) sees pixel (2k+py, 2j+px) at tap (py - 2a + 1, px - 2b + 1); taps outside 0..2 do not exist
#pragma unroll
        for (int a = 0; a <= py; ++a)
#pragma unroll
          for (int b = 0; b <= px; ++b) {
            const uint32_t me = static_cast<uint32_t>((py - 2 * a + 1) * 3 + (px - 2 * b + 1));
            const uint32_t hit_lo = __vcmpeq4(rec[a][b].x, me * 0x01010101u), hit_hi = __vcmpeq4(rec[a][b].y, me * 0x01010101u);
            const uint32_t gw[4] = {g[a][b].x, g[a][b].y, g[a][b].z, g[a][b].w};
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const uint32_t hit = ((e < 4 ? hit_lo : hit_hi) >> (8 * (e & 3))) & 1u;
              const float f = __uint_as_float((e & 1) ? (gw[e >> 1] & 0xffff0000u) : (gw[e >> 1] << 16));
              acc[e] += hit ? f : 0.f;
            }
          }
        const long long xoff = ((1LL * n * H + iy) * W + ix) * C + c8 * 8;
        if (x != nullptr) {
          const uint4 xv = *reinterpret_cast<const uint4*>(x + xoff);
          const __nv_bfloat162* x2 = reinterpret_cast<const __nv_bfloat162*>(&xv);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = __bfloat1622float2(x2[e]);
            if (!(f.x > 0.f)) acc[2 * e] = 0.f;
            if (!(f.y > 0.f)) acc[2 * e + 1] = 0.f;
          }
        }
        uint4 o;
        __nv_bfloat162* o2 = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
        for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(acc[2 * e], acc[2 * e + 1]);
        *reinterpret_cast<uint4*>(dx + xoff) = o;
      }
    }
  }
}
int maxpool_bwd(const bf16* dy, const uint8_t* idx, const bf16* x, bf16* dx, int N, int H, int W, int C, cudaStream_t st) {
  ACT_CHECK(C % 8 == 0, "maxpool needs C % 8 == 0");
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = 1LL * N * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16));
  ACT_CUDA(launch_pdl(maxpool_bwd_kernel, dim3(blocks), dim3(256), 0, st, dy, idx, x, dx, N, H, W, C, Ho, Wo));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ stem image pack
// policy.py:268-272 normalisation fused with the layout change the implicit-GEMM stem conv reads: pixel x of row y
// lands at out[b][y][x + 3][0..3] = {R, G, B, 0} (bf16), columns [0,3) and [W+3, Wp) are zero (horizontal padding).
// U8 = the dataset's native layout (utils.py:133-158 keeps uint8 HWC frames and divides by 255 on the host): pixel
// (y, x) of camera `cam` is img[b][cam][y][x][0..3) uint8; x/255 is formed in fp32 exactly like `image_data / 255.0`.
// One thread = four consecutive output pixels (two 16-byte stores); interior groups of the uint8 form read their twelve
// bytes as four aligned words re-aligned with funnel shifts (the one-pixel-per-thread form ran at 2.2 TB/s).
template <bool U8>
__global__ void __launch_bounds__(256) stem_pack_image_kernel(const void* __restrict__ img_, long long img_bs, bf16* __restrict__ out,
                                                              int B, int H, int W, int Wp) {
  pdl_trigger(); pdl_wait();
  const float mean[3] = {0.485f, 0.456f, 0.406f};
  const float istd[3] = {1.f / 0.229f, 1.f / 0.224f, 1.f / 0.225f};
  const int G = (Wp + 3) / 4;
  const long long total = 1LL * B * H * G;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    const int q = static_cast<int>(i % G);
    const long long r = i / G;
    const int y = static_cast<int>(r % H), b = static_cast<int>(r / H);
    const int x0 = 4 * q - 3;
    float v[4][3];
#pragma unroll
    for (int j = 0; j < 4; ++j) { v[j][0] = 0.f; v[j][1] = 0.f; v[j][2] = 0.f; }
    const bool interior = x0 >= 0 && x0 + 5 < W;
    if (U8) {
      const uint8_t* px = static_cast<const uint8_t*>(img_) + b * img_bs + (1LL * y * W + x0) * 3;
      if (interior) {
        const uint32_t mis = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(px) & 3u);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(px - mis);
        const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2), w3 = __ldg(w + 3);
        const uint32_t sh = mis * 8u;
        const uint32_t r3[3] = {__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh)};
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int k = 3 * j + c;
            v[j][c] = (static_cast<float>((r3[k >> 2] >> (8 * (k & 3))) & 0xffu) / 255.0f - mean[c]) * istd[c];
          }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (x0 + j >= 0 && x0 + j < W) {
#pragma unroll
            for (int c = 0; c < 3; ++c) v[j][c] = (static_cast<float>(__ldg(px + 3 * j + c)) / 255.0f - mean[c]) * istd[c];
          }
      }
    } else {
      const float* im = static_cast<const float*>(img_) + b * img_bs + 1LL * y * W + x0;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (interior || (x0 + j >= 0 && x0 + j < W)) {
#pragma unroll
          for (int c = 0; c < 3; ++c) v[j][c] = (__ldg(im + 1LL * c * H * W + j) - mean[c]) * istd[c];
        }
    }
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v[j][0], v[j][1]), hi = __floats2bfloat162_rn(v[j][2], 0.f);
      o[2 * j] = *reinterpret_cast<uint32_t*>(&lo); o[2 * j + 1] = *reinterpret_cast<uint32_t*>(&hi);
    }
    uint2* dst = reinterpret_cast<uint2*>(out) + (1LL * b * H + y) * Wp + 4 * q;     // 8 bytes per pixel
    if (4 * q + 3 < Wp && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
      reinterpret_cast<uint4*>(dst)[0] = make_uint4(o[0], o[1], o[2], o[3]);
      reinterpret_cast<uint4*>(dst)[1] = make_uint4(o[4], o[5], o[6], o[7]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (4 * q + j < Wp) dst[j] = make_uint2(o[2 * j], o[2 * j + 1]);
    }
  }
}
int stem_pack_image_u8(const uint8_t* img_cam, long long img_bs, bf16* out, int B, int H, int W, int Wp, cudaStream_t st) {
  const long long total = 1LL * B * H * ((Wp + 3) / 4);
  stem_pack_image_kernel<true><<<static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16)), 256, 0, st>>>(img_cam, img_bs, out, B, H, W, Wp);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
int stem_pack_image(const float* img_cam, long long img_bs, bf16* out, int B, int H, int W, int Wp, cudaStream_t st) {
  const long long total = 1LL * B * H * ((Wp + 3) / 4);
  stem_pack_image_kernel<false><<<static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 16)), 256, 0, st>>>(img_cam, img_bs, out, B, H, W, Wp);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ dropout utilities
__global__ void dropout_mask_kernel(DropDesc d, uint32_t step, long long n8, long long n, uint8_t* __restrict__ out) {
  for (long long g = blockIdx.x * 1LL * blockDim.x + threadIdx.x; g < n8; g += gridDim.x * 1LL * blockDim.x) {
    const uint32_t m = drop_keep8(d, static_cast<unsigned long long>(g), step);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (8 * g + i < n) out[8 * g + i] = (m >> i) & 1u;
  }
}
uint32_t dropout_threshold(float p) {
  const double keep = 1.0 - static_cast<double>(p);
  long long t = static_cast<long long>(keep * 65536.0 + 0.5);
  return static_cast<uint32_t>(std::min<long long>(std::max<long long>(t, 0), 0x10000));
}
int dropout_mask(unsigned long long seed, uint32_t site, uint32_t step, float p, long long n, uint8_t* out, cudaStream_t st) {
  ACT_CHECK(p >= 0.f && p < 1.f && n >= 0 && out != nullptr, "dropout_mask: bad arguments");
  if (n == 0) return 0;
  DropDesc d;
  d.seed_lo = static_cast<uint32_t>(seed); d.seed_hi = static_cast<uint32_t>(seed >> 32); d.site = site;
  d.thresh = dropout_threshold(p); d.scale = 1.f / (1.f - p);
  const long long n8 = (n + 7) / 8;
  dropout_mask_kernel<<<static_cast<int>(std::min<long long>((n8 + 255) / 256, 148 * 8)), 256, 0, st>>>(d, step, n8, n, out);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
__global__ void bump_counter_kernel(uint32_t* c) { *c += 1u; }
int bump_counter(uint32_t* counter, cudaStream_t st) {
  bump_counter_kernel<<<1, 1, 0, st>>>(counter);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace act
