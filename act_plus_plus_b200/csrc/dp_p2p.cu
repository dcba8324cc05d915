// Data-parallel optimiser step fused with its gradient exchange over NVLink peer memory (the reference is
// single-GPU: imitate_episodes.py:603-607 is loss.backward(); optimizer.step()).
//
// Every rank maps every other rank's flat gradient / parameter / bf16-shadow buffers (CUDA IPC, dp_export / dp_connect
// in engine_params.cu).  For one contiguous gradient range the three steps "all-reduce the gradients, run AdamW on the
// whole range, (implicitly) keep replicas identical" become ONE kernel per rank:
//   rank r owns the r-th slice of the range; for each float4 of its slice it loads the gradient from all ranks (peer
//   loads, fixed rank order -> run-to-run deterministic), averages, applies torch.optim.AdamW (detr/main.py:102-110)
//   with its private exp_avg / exp_avg_sq slice, and stores the new fp32 parameter and bf16 shadow into EVERY rank's
//   buffers (peer stores).  Reduce-scatter, optimiser and all-gather in one pass: each rank reads 4 B/param from each
//   peer for 1/N of the range and writes 4+2 B/param to each peer, instead of a ring all-reduce (2x the gradient bytes
//   through local HBM on top of NVLink) followed by a full-range AdamW pass (28 B/param of local HBM per rank).
// Ordering between ranks is a flag barrier in peer memory (one slot per barrier site and rank, monotonically increasing
// epochs kept on the device so a captured CUDA graph can replay it).
#include "kernels.h"
#include <cmath>

namespace act {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
// L2-only (".cg") load: the data was published by the owner's release store before the barrier this kernel is ordered
// after, and L1 is bypassed, so no stale line can be hit.  (System-scope relaxed loads were measured 4-5x slower here.)
__device__ __forceinline__ float4 ld_cg_v4(const float4* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
// NVLS (NVSwitch multicast) accesses on a multicast address that is bound to the SAME offset of every rank's buffer:
// one load returns the switch-reduced sum over all ranks, one store lands in every rank's memory.  The link carries the
// payload once per rank instead of once per (rank, peer) pair.
__device__ __forceinline__ float4 multimem_ld_reduce_v4(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(float4* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_b64(uint2* p, const uint2& v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// One warp.  Lane j < world: publish "rank `rank` reached site `site` for the e-th time" in rank j's signal block, then
// wait until rank j has published the same in ours.  status[0] is set to 1 if a peer does not show up within
// `timeout_ns` (the step's results are then meaningless, but the GPU is not left spinning for ever).
__global__ void dp_barrier_kernel(DpPeers peers, uint32_t* __restrict__ epoch, uint32_t* __restrict__ status, int site,
                                  unsigned long long timeout_ns) {
  const int lane = threadIdx.x;
  const uint32_t e = epoch[site] + 1;
  // an earlier barrier of this step timed out: do not wait again (the exchange kernels that follow skip their work,
  // so no replica is overwritten with gradients a peer never finished); the host raises when it polls the status
  if (status[0] != 0u) { if (lane == 0) epoch[site] = e; return; }
  __syncwarp();
  __threadfence_system();                                    // everything this GPU wrote before the barrier
  if (lane < peers.world) {
    st_release_sys(peers.signal[lane] + site * DP_MAX_WORLD + peers.rank, e);
    const uint32_t* mine = peers.signal[peers.rank] + site * DP_MAX_WORLD + lane;
    const unsigned long long t0 = globaltimer_ns();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - e) < 0) {
      __nanosleep(200);
      if (globaltimer_ns() - t0 > timeout_ns) { status[0] = 1u; break; }
    }
  }
  __syncwarp();
  __threadfence_system();
  if (lane == 0) epoch[site] = e;
}

int dp_barrier(const DpPeers& peers, uint32_t* epoch, uint32_t* status, int site, cudaStream_t st) {
  ACT_CHECK(site >= 0 && site < DP_MAX_SITES, "dp_barrier: too many barrier sites in one step");
  dp_barrier_kernel<<<1, 32, 0, st>>>(peers, epoch, status, site, 20ull * 1000ull * 1000ull * 1000ull);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// Rank r's slice of n4 float4s: [r*n4/world, (r+1)*n4/world) with the remainder spread over the lowest ranks.
__host__ __device__ inline void dp_slice(long long n4, int rank, int world, long long* lo, long long* hi) {
  const long long base = n4 / world, rem = n4 % world;
  *lo = rank * base + (rank < rem ? rank : rem);
  *hi = *lo + base + (rank < rem ? 1 : 0);
}

template <int WORLD, int U>
__global__ void __launch_bounds__(256) dp_adamw_kernel(DpPeers peers, AdamWArgs a, long long off, float bc1_host, float rs_host,
                                                       int mode) {
  if (a.abort_flag != nullptr && *a.abort_flag != 0u) return;    // a barrier of this step failed: leave every replica as it is
  float bc1, rsqrt_bc2;
  adam_bias(a, bc1_host, rs_host, &bc1, &rsqrt_bc2);
  // WORLD > 0: compile-time rank count; WORLD == 0: run-time count.  Each thread owns U float4 chunks per iteration and
  // the loop is software-pipelined over two register buffers: the U * world peer gradient loads and the local p / m / v
  // loads of iteration k+1 are issued BEFORE iteration k is computed and stored, so a remote load (~2 us) always has a
  // full iteration of stores and math to hide behind (bytes in flight set the NVLink rate).
  // mode bits (experiments, tools/dp_exchange_time.py): 1 = system-scope loads, 2 = read the local gradient W times
  // instead of the peers' (what the loads cost without NVLink), 4 = store to the local replica only.
  constexpr bool MC = WORLD < 0;                               // WORLD == -1: NVLS multicast, any rank count
  constexpr int W = MC ? 1 : WORLD > 0 ? WORLD : DP_MAX_WORLD;
  const int world = WORLD > 0 ? WORLD : peers.world;
  const long long n4 = a.n / 4;
  long long lo, hi;
  dp_slice(n4, peers.rank, world, &lo, &hi);
  const float decay = 1.f - a.lr * a.wd;
  const float step_size = a.lr / bc1;
  const float inv_world = 1.f / static_cast<float>(world);
  const long long off4 = off / 4;                              // ranges are 16-byte aligned (checked on the host)
  const long long S = gridDim.x * 1LL * blockDim.x;
  const bool ld_sys = mode & 1, ld_local = mode & 2, st_local = mode & 4;
  struct Buf { float4 g[U][W]; float4 p[U], m[U], v[U]; };
  auto issue = [&](Buf& b, long long i0) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * S;
      if (i < hi) {
#pragma unroll
        for (int j = 0; j < W; ++j)
          if (j < (MC ? 1 : world)) {
            if (MC) {
              b.g[u][0] = multimem_ld_reduce_v4(reinterpret_cast<const float4*>(peers.mc_grads) + off4 + i);
            } else {
              const float4* src = reinterpret_cast<const float4*>(peers.grads[ld_local ? peers.rank : j]) + off4 + i;
              b.g[u][j] = ld_sys ? ld_relaxed_sys_v4(src) : ld_cg_v4(src);
            }
          }
        b.p[u] = reinterpret_cast<const float4*>(a.p)[i];
        b.m[u] = reinterpret_cast<const float4*>(a.m)[i];
        b.v[u] = reinterpret_cast<const float4*>(a.v)[i];
      }
    }
  };
  auto finish = [&](Buf& b, long long i0) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * S;
      if (i >= hi) continue;
      float4 g = b.g[u][0];
#pragma unroll
      for (int j = 1; j < W; ++j)
        if (!MC && j < world) { g.x += b.g[u][j].x; g.y += b.g[u][j].y; g.z += b.g[u][j].z; g.w += b.g[u][j].w; }
      g.x *= inv_world; g.y *= inv_world; g.z *= inv_world; g.w *= inv_world;
      float4 p = b.p[u], m = b.m[u], v = b.v[u];
#define ADAM1(c)                                                  \
      p.c *= decay;                                               \
      m.c = a.beta1 * m.c + (1.f - a.beta1) * g.c;                \
      v.c = a.beta2 * v.c + (1.f - a.beta2) * g.c * g.c;          \
      p.c -= step_size * (m.c / (sqrtf(v.c) * rsqrt_bc2 + a.eps));
      ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
      reinterpret_cast<float4*>(a.m)[i] = m;
      reinterpret_cast<float4*>(a.v)[i] = v;
      uint2 o;
      {
        __nv_bfloat162 l2 = __floats2bfloat162_rn(p.x, p.y), h2 = __floats2bfloat162_rn(p.z, p.w);
        o.x = *reinterpret_cast<uint32_t*>(&l2); o.y = *reinterpret_cast<uint32_t*>(&h2);
      }
      if (MC) {
        multimem_st_v4(reinterpret_cast<float4*>(peers.mc_params) + off4 + i, p);
        if (a.shadow) multimem_st_b64(reinterpret_cast<uint2*>(peers.mc_shadow) + off4 + i, o);
      } else {
#pragma unroll
        for (int j = 0; j < W; ++j) {
          if (j < world && (!st_local || j == peers.rank)) {
            reinterpret_cast<float4*>(peers.params[j])[off4 + i] = p;
            if (a.shadow) reinterpret_cast<uint2*>(peers.shadow[j])[off4 + i] = o;
          }
        }
      }
    }
  };
  {
    Buf b0, b1;
    long long i0 = lo + blockIdx.x * 1LL * blockDim.x + threadIdx.x;
    const long long stride = U * S;
    issue(b0, i0);
    while (i0 < hi) {
      issue(b1, i0 + stride);
      finish(b0, i0);
      i0 += stride;
      if (i0 >= hi) break;
      issue(b0, i0 + stride);
      finish(b1, i0);
      i0 += stride;
    }
  }
  // tail (n % 4 elements): rank 0, first threads of block 0
  if (peers.rank == 0 && blockIdx.x == 0 && threadIdx.x < (a.n & 3)) {
    const long long i = n4 * 4 + threadIdx.x;
    float g = 0.f;
    for (int j = 0; j < world; ++j) g += peers.grads[j][off + i];
    g *= inv_world;
    float p = a.p[i] * decay;
    const float m = a.beta1 * a.m[i] + (1.f - a.beta1) * g;
    const float v = a.beta2 * a.v[i] + (1.f - a.beta2) * g * g;
    a.m[i] = m; a.v[i] = v;
    p -= step_size * (m / (sqrtf(v) * rsqrt_bc2 + a.eps));
    for (int j = 0; j < world; ++j) {
      peers.params[j][off + i] = p;
      if (a.shadow) peers.shadow[j][off + i] = __float2bfloat16_rn(p);
    }
  }
  // Peer stores must be performed before the next barrier's flag.  One cumulative system fence per block after a block
  // barrier (a fence in every thread was measured to serialise: ~3 us of NVLink round trip per warp, 110 us per launch).
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
}

__global__ void multimem_reduce_kernel(const float4* __restrict__ mc, float4* __restrict__ dst, long long n4) {
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * 1LL * blockDim.x)
    dst[i] = multimem_ld_reduce_v4(mc + i);
}
__global__ void multimem_bcast_kernel(float4* __restrict__ mc, const float4* __restrict__ src, long long n4) {
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * 1LL * blockDim.x)
    multimem_st_v4(mc + i, src[i]);
  __syncthreads();
  if (threadIdx.x == 0) __threadfence_system();
}
int multimem_reduce_f32(const float* mc_src, float* dst_local, long long n, cudaStream_t st) {
  ACT_CHECK(mc_src && dst_local && n > 0 && (n & 3) == 0, "multimem_reduce_f32: n must be a positive multiple of 4");
  ACT_CHECK(((reinterpret_cast<uintptr_t>(mc_src) | reinterpret_cast<uintptr_t>(dst_local)) & 15) == 0, "multimem_reduce_f32: 16-byte alignment");
  multimem_reduce_kernel<<<static_cast<int>(std::min<long long>(148 * 8, (n / 4 + 255) / 256)), 256, 0, st>>>(
      reinterpret_cast<const float4*>(mc_src), reinterpret_cast<float4*>(dst_local), n / 4);
  ACT_CUDA(cudaGetLastError());
  return 0;
}
int multimem_bcast_f32(float* mc_dst, const float* src_local, long long n, cudaStream_t st) {
  ACT_CHECK(mc_dst && src_local && n > 0 && (n & 3) == 0, "multimem_bcast_f32: n must be a positive multiple of 4");
  ACT_CHECK(((reinterpret_cast<uintptr_t>(mc_dst) | reinterpret_cast<uintptr_t>(src_local)) & 15) == 0, "multimem_bcast_f32: 16-byte alignment");
  multimem_bcast_kernel<<<static_cast<int>(std::min<long long>(148 * 8, (n / 4 + 255) / 256)), 256, 0, st>>>(
      reinterpret_cast<float4*>(mc_dst), reinterpret_cast<const float4*>(src_local), n / 4);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// a.p / a.m / a.v point at this rank's own buffers at the START of the range (offset `off` floats into the flat buffers);
// a.g and a.shadow's value are ignored except that a.shadow != nullptr asks for the bf16 shadow to be written.
int dp_adamw_step(const DpPeers& peers, const AdamWArgs& a, long long off, cudaStream_t st) {
  if (a.n <= 0) return 0;
  ACT_CHECK(peers.world >= 2 && peers.world <= DP_MAX_WORLD, "dp_adamw_step: peers not connected");
  ACT_CHECK((off & 3) == 0 && (reinterpret_cast<uintptr_t>(a.p) & 15) == 0, "data-parallel AdamW range must be 16-byte aligned");
  const double bc1 = 1.0 - std::pow(static_cast<double>(a.beta1), a.step);
  const double bc2 = 1.0 - std::pow(static_cast<double>(a.beta2), a.step);
  const long long n4 = a.n / 4, slice = (n4 + peers.world - 1) / peers.world;
  static int mode = -1, per_sm_env = -1;
  if (mode < 0) { const char* e = getenv("ACT_DP_LD_SYS"); mode = e ? atoi(e) & 7 : 0; }
  if (per_sm_env < 0) { const char* e = getenv("ACT_DP_PERSM"); per_sm_env = e ? atoi(e) : 0; }
  const bool mc = peers.mc_grads != nullptr && peers.mc_params != nullptr && (a.shadow == nullptr || peers.mc_shadow != nullptr);
  const int unroll = mc ? 4 : peers.world == 2 ? 2 : 1;
  const float f1 = static_cast<float>(bc1), f2 = static_cast<float>(1.0 / std::sqrt(bc2));
  auto launch = [&](auto kern) -> int {
    static int occ_tab[DP_MAX_WORLD + 1] = {};                   // resident 256-thread blocks per SM of each instantiation
    int& occ = occ_tab[mc ? 0 : peers.world];
    if (occ == 0) { ACT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, 0)); occ = std::max(1, occ); }
    int per_sm = a.background > 0 ? std::min(a.background, occ) : occ;
    if (per_sm_env > 0) per_sm = std::min(per_sm_env, occ);
    const int blocks = static_cast<int>(std::max<long long>(1, std::min<long long>((slice + 256LL * unroll - 1) / (256LL * unroll), 148LL * per_sm)));
    kern<<<blocks, 256, 0, st>>>(peers, a, off, f1, f2, mode);
    return 0;
  };
  int rc;
  switch (mc ? -1 : peers.world) {
    case -1: rc = launch(dp_adamw_kernel<-1, 4>); break;
    case 2: rc = launch(dp_adamw_kernel<2, 2>); break;
    case 4: rc = launch(dp_adamw_kernel<4, 1>); break;
    case 8: rc = launch(dp_adamw_kernel<8, 1>); break;
    default: rc = launch(dp_adamw_kernel<0, 1>); break;
  }
  if (rc) return rc;
  ACT_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace act
