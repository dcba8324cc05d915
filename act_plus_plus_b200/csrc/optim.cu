// Fused AdamW (torch.optim.AdamW semantics, detr/main.py:102-110) and the temporal-ensemble step
// (imitate_episodes.py:403-411).  Both are HBM-bound: 128-bit accesses, grid sized to the SM count.
#include "kernels.h"
#include <cmath>

namespace act {

// p *= 1 - lr*wd; m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
// Also refreshes the bf16 shadow used as tensor-core operand.  mode 1: gradient identically zero (dead decoder
// layers) -> m, v stay 0 and only the decoupled decay touches p.
__global__ void adamw_kernel(AdamWArgs a, float bc1_host, float rs_host) {
  pdl_trigger(); pdl_wait();
  float bc1, rsqrt_bc2;
  adam_bias(a, bc1_host, rs_host, &bc1, &rsqrt_bc2);
  const long long n4 = a.n / 4;
  const float decay = 1.f - a.lr * a.wd;
  const float step_size = a.lr / bc1;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * 1LL * blockDim.x) {
    float4 p = reinterpret_cast<float4*>(a.p)[i];
    if (a.mode == 0) {
      const float4 g = reinterpret_cast<const float4*>(a.g)[i];
      float4 m = reinterpret_cast<float4*>(a.m)[i], v = reinterpret_cast<float4*>(a.v)[i];
#define ADAM1(c)                                                      \
      p.c *= decay;                                                   \
      m.c = a.beta1 * m.c + (1.f - a.beta1) * g.c;                    \
      v.c = a.beta2 * v.c + (1.f - a.beta2) * g.c * g.c;              \
      p.c -= step_size * (m.c / (sqrtf(v.c) * rsqrt_bc2 + a.eps));
      ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
      reinterpret_cast<float4*>(a.m)[i] = m;
      reinterpret_cast<float4*>(a.v)[i] = v;
    } else {
      p.x *= decay; p.y *= decay; p.z *= decay; p.w *= decay;
    }
    reinterpret_cast<float4*>(a.p)[i] = p;
    if (a.shadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(p.x, p.y), hi = __floats2bfloat162_rn(p.z, p.w);
      uint2 o;
      o.x = *reinterpret_cast<uint32_t*>(&lo); o.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(a.shadow)[i] = o;
    }
  }
  // tail (n % 4) handled by the first threads of block 0
  if (blockIdx.x == 0 && threadIdx.x < (a.n & 3)) {
    const long long i = n4 * 4 + threadIdx.x;
    float p = a.p[i] * decay;
    if (a.mode == 0) {
      const float g = a.g[i];
      const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * g;
      const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * g * g;
      a.m[i] = m; a.v[i] = v;
      p -= step_size * (m / (sqrtf(v) * rsqrt_bc2 + a.eps));
    }
    a.p[i] = p;
    if (a.shadow) a.shadow[i] = __float2bfloat16_rn(p);
  }
}

int adamw_step(const AdamWArgs& a, cudaStream_t st) {
  if (a.n <= 0) return 0;
  ACT_CHECK((reinterpret_cast<uintptr_t>(a.p) & 15) == 0, "AdamW segment must be 16-byte aligned");
  const double bc1 = 1.0 - std::pow(static_cast<double>(a.beta1), a.step);
  const double bc2 = 1.0 - std::pow(static_cast<double>(a.beta2), a.step);
  const long long n4 = a.n / 4;
  const int blocks = static_cast<int>(std::max<long long>(1, std::min<long long>((n4 + 255) / 256, 148 * (a.background > 0 ? std::min(a.background, 8) : 8))));
  ACT_CUDA(launch_pdl(adamw_kernel, dim3(blocks), dim3(256), 0, st, a, static_cast<float>(bc1), static_cast<float>(1.0 / std::sqrt(bc2))));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

__global__ void set_counter_kernel(int* c, int value, int increment) { *c = increment ? *c + 1 : value; }
int set_counter(int* counter, int value, bool increment, cudaStream_t st) {
  set_counter_kernel<<<1, 1, 0, st>>>(counter, value, increment ? 1 : 0);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------ temporal ensemble
// Ring layout of all_time_actions: the chunk predicted at time t' lives in ring[e][t' % nq][0..nq)[A]; entry
// (t', t) of the reference's [T, T+nq, A] buffer is ring[e][t' % nq][t - t'].  One block per environment:
//   1. store the new chunk (row t), 2. for t' = max(0,t-nq+1)..t in ascending order keep rows whose A values are all
//   non-zero (imitate_episodes.py:405), 3. weights exp(-k*i)/sum over the kept rows in that order (float64),
//   4. out[e][a] = sum_i float64(row_i[a]) * w_i.
__global__ void temporal_agg_kernel(float* __restrict__ ring, const float* __restrict__ chunks, const int* __restrict__ tptr,
                                    double* __restrict__ out, int nq, int A, double k, const float* __restrict__ amean,
                                    const float* __restrict__ astd) {
  extern __shared__ unsigned char sm[];
  int* keep = reinterpret_cast<int*>(sm);                 // [nq] rank among kept rows or -1
  __shared__ int n_keep;
  const int e = blockIdx.x, tid = threadIdx.x;
  const int t = tptr[e];
  float* R = ring + 1LL * e * nq * nq * A;
  const float* c = chunks + 1LL * e * nq * A;
  float* slot = R + 1LL * (t % nq) * nq * A;
  for (int i = tid; i < nq * A; i += blockDim.x) slot[i] = c[i];
  __syncthreads();
  const int t0 = max(0, t - nq + 1);
  const int n = t - t0 + 1;
  for (int i = tid; i < n; i += blockDim.x) {
    const int tp = t0 + i;
    const float* row = R + (1LL * (tp % nq) * nq + (t - tp)) * A;
    int ok = 1;
    for (int a = 0; a < A; ++a) ok &= (row[a] != 0.f);
    keep[i] = ok;
  }
  __syncthreads();
  if (tid == 0) {
    int cnt = 0;
    for (int i = 0; i < n; ++i) { if (keep[i]) keep[i] = cnt++; else keep[i] = -1; }
    n_keep = cnt;
  }
  __syncthreads();
  if (tid < A) {
    double wsum = 0.0;
    for (int i = 0; i < n_keep; ++i) wsum += exp(-k * static_cast<double>(i));
    double acc = 0.0;
    for (int i = 0; i < n; ++i) {
      if (keep[i] < 0) continue;
      const int tp = t0 + i;
      const float v = R[(1LL * (tp % nq) * nq + (t - tp)) * A + tid];
      acc += static_cast<double>(v) * (exp(-k * static_cast<double>(keep[i])) / wsum);
    }
    // post_process (imitate_episodes.py:294): float64 raw action * float32 std + float32 mean, promoted to float64
    if (astd) acc = acc * static_cast<double>(astd[tid]) + static_cast<double>(amean[tid]);
    out[1LL * e * A + tid] = acc;
  }
}

__global__ void chunk_select_kernel(const float* __restrict__ chunks, const int* __restrict__ tptr, double* __restrict__ out,
                                    int E, int nq, int A, int base_delay, const float* __restrict__ amean,
                                    const float* __restrict__ astd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E * A) return;
  const int e = i / A, a = i % A;
  const int j = tptr[e] % (nq - base_delay) + ((base_delay > 0 && a >= A - 2) ? base_delay : 0);
  double v = static_cast<double>(chunks[(1LL * e * nq + j) * A + a]);
  if (astd) v = v * static_cast<double>(astd[a]) + static_cast<double>(amean[a]);
  out[i] = v;
}
int chunk_select(const float* chunks, const int* t, double* out, int E, int nq, int A, int base_delay,
                 const float* action_mean, const float* action_std, cudaStream_t st) {
  ACT_CHECK(base_delay >= 0 && base_delay < nq && A >= 2 && (!action_std == !action_mean), "chunk_select: bad arguments");
  if (E <= 0) return 0;
  chunk_select_kernel<<<(E * A + 127) / 128, 128, 0, st>>>(chunks, t, out, E, nq, A, base_delay, action_mean, action_std);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

int temporal_agg_step(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A, double k,
                      cudaStream_t st, const float* action_mean, const float* action_std) {
  ACT_CHECK(A <= 128 && nq >= 1 && (!action_std == !action_mean), "temporal_agg: action_dim <= 128, mean and std together");
  if (E <= 0) return 0;
  temporal_agg_kernel<<<E, 128, nq * sizeof(int), st>>>(ring, chunks, t, out, nq, A, k, action_mean, action_std);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace act
