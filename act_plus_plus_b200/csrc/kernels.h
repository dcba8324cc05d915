// Host-callable launchers of the bandwidth-bound kernels (pack / normalise / softmax / loss / optimizer ...).
#pragma once
#include <algorithm>
#include "tc_api.h"

namespace act {

// ---- pack.cu
int cast_bf16(const float* src, bf16* dst, long long n, cudaStream_t st);
int pack_conv_weights(const float* w, const float* scale, bf16* wpk, bf16* wpkt, int Cout, int Cin, int KK, cudaStream_t st);
struct PackJob { const float* w; const float* scale; bf16* wpk; bf16* wpkt; int Cout, Cin, KK, tile0; };
struct PackJobs { static constexpr int MAX = 64; int n = 0; PackJob job[MAX]; };
int pack_conv_batched(PackJobs& jobs, cudaStream_t st);   // all jobs in one launch
struct PairJob { const bf16* src; bf16* dst; int8_t even_slot[18]; int8_t odd_slot[18]; };
struct PairJobs { static constexpr int MAX = 24; int n = 0; PairJob job[MAX]; };
int pack_pair_batched(const PairJobs& jobs, cudaStream_t st);   // paired 64-channel 3x3 operands, all jobs in one launch
int bn_fold(const float* w, const float* b, const float* rm, const float* rv, float* scale, float* shift, int n, cudaStream_t st);
int pack_stem_weights(const float* w, bf16* wpk, cudaStream_t st);              // [64,3,7,7] fp32 -> [64][7*64] bf16 (row-window layout)
int pack_conv1_weights(const float* w, bf16* wpk, int KP, cudaStream_t st);     // [64,147] fp32 -> [64,KP] bf16, zero padded

// ---- fused_norm.cu
int ln_fwd(const float* t, const float* gamma, const float* beta, const float* pos, int L, float* y, bf16* yb, bf16* ypb,
           float* mean, float* rstd, int M, int D, cudaStream_t st);
int ln_bwd(const float* dy, const float* dy2, const float* t, const float* mean, const float* rstd, const float* gamma,
           float* dt, bf16* dtb, float* dgamma, float* dbeta, int M, int D, cudaStream_t st, const DropDesc* drop = nullptr);
// uint8 keep mask (1 = kept) of elements [0, n) of dropout stream (seed, site, step): what the fused kernels regenerate
int dropout_mask(unsigned long long seed, uint32_t site, uint32_t step, float p, long long n, uint8_t* out, cudaStream_t st);
int bump_counter(uint32_t* counter, cudaStream_t st);
uint32_t dropout_threshold(float p);        // keep iff 16-bit lane < threshold = round((1 - p) * 65536)
int add_pos_cast(const float* x, const float* pos, int L, bf16* xb, bf16* xpb, int M, int D, cudaStream_t st);
int softmax_fwd(const float* S, const uint8_t* mask, bf16* P, int R, int L, int Lp, int rows_per_batch, cudaStream_t st);
int softmax_bwd(const bf16* P, const float* dP, bf16* dS, int R, int L, int Lp, float scale, cudaStream_t st);
int colsum_bf16(const bf16* X, long long ld, int M, int N, float* out, cudaStream_t st);

// ---- fused_misc.cu
struct IdxMap {            // off(idx) = (idx / rpb) * bs + (idx % rpb) * s
  int rpb = 1 << 30;
  long long bs = 0, s = 0;
};
inline IdxMap lin(long long s) { IdxMap m; m.s = s; return m; }
inline IdxMap batched(int rpb, long long bs, long long s) { IdxMap m; m.rpb = rpb; m.bs = bs; m.s = s; return m; }
struct SgemmArgs {         // C(i,j) (+)= alpha * sum_r A(i,r) * B(r,j) (+ bias[j])
  int I = 0, J = 0, R = 0;
  const float* A = nullptr; IdxMap ai, ar;
  const float* B = nullptr; IdxMap br, bj;
  float* C = nullptr; IdxMap ci, cj;
  const float* bias = nullptr;
  float alpha = 1.f;
  int accumulate = 0;
  // optional z-score of the A operand along r (rollout pre-processing, imitate_episodes.py:290): (A - a_mean[r]) / a_std[r]
  const float* a_mean = nullptr; const float* a_std = nullptr;
};
int sgemm_generic(const SgemmArgs& a, cudaStream_t st);
int act_loss(const float* a_hat, const float* actions, long long act_bs, const uint8_t* is_pad, long long pad_bs,
             const float* mu, const float* logvar, int B, int nq, int A, int Z, float kl_weight, float gscale,
             float* losses, float* d_a_hat, float* d_mu, float* d_logvar, cudaStream_t st, const float* gvec = nullptr);
int reparam_fwd(const float* info, const float* eps, float* mu, float* logvar, float* z, int B, int Z, cudaStream_t st);
int reparam_bwd(const float* dz, const float* eps, const float* logvar, const float* d_mu_kl, const float* d_lv_kl,
                float* d_info, int B, int Z, cudaStream_t st);
int maxpool_fwd(const bf16* x, bf16* y, uint8_t* idx, int N, int H, int W, int C, cudaStream_t st);
int maxpool_bwd(const bf16* dy, const uint8_t* idx, const bf16* x, bf16* dx, int N, int H, int W, int C, cudaStream_t st);
// image fp32 [B,3,H,W] (batch stride img_bs) -> ImageNet-normalised NHWC4 bf16 [B][H][Wp][4], 3 zero pixels left pad
int stem_pack_image(const float* img_cam, long long img_bs, bf16* out, int B, int H, int W, int Wp, cudaStream_t st);
int stem_pack_image_u8(const uint8_t* img_cam, long long img_bs, bf16* out, int B, int H, int W, int Wp, cudaStream_t st);

// ---- optim.cu
struct AdamWArgs {
  float* p; const float* g; float* m; float* v; bf16* shadow;   // shadow may be null
  long long n;
  float lr, beta1, beta2, eps, wd;
  int step;          // 1-based
  const int* step_dev = nullptr;   // when set, the kernels read the step from this device counter instead (bias corrections
                                   // computed on the device): a captured CUDA graph then advances with every replay
  const uint32_t* abort_flag = nullptr;   // data-parallel kernels: non-zero = an earlier barrier failed, do nothing
  int mode;          // 0 = full AdamW, 1 = weight decay only (gradient identically zero)
  int background = 0;   // n > 0: launch n blocks per SM (n x 256 threads) so concurrently running tensor-core CTAs still find a slot
};
int adamw_step(const AdamWArgs& a, cudaStream_t st);
int set_counter(int* counter, int value, bool increment, cudaStream_t st);      // counter = value, or counter += 1
// bias corrections of torch.optim.AdamW for step t: bc1 = 1 - b1^t, rs = 1 / sqrt(1 - b2^t); block-uniform
__device__ __forceinline__ void adam_bias(const AdamWArgs& a, float bc1_host, float rs_host, float* bc1, float* rs) {
  if (a.step_dev == nullptr) { *bc1 = bc1_host; *rs = rs_host; return; }
  __shared__ float s_bias[2];
  if (threadIdx.x == 0) {
    const double t = static_cast<double>(*a.step_dev);
    s_bias[0] = static_cast<float>(1.0 - pow(static_cast<double>(a.beta1), t));
    s_bias[1] = static_cast<float>(1.0 / sqrt(1.0 - pow(static_cast<double>(a.beta2), t)));
  }
  __syncthreads();
  *bc1 = s_bias[0]; *rs = s_bias[1];
}

// ---- dp_p2p.cu: data-parallel AdamW fused with the gradient exchange over NVLink peer memory
constexpr int DP_MAX_WORLD = 8, DP_MAX_SITES = 64;
struct DpPeers {           // device pointers valid on THIS device: index `rank` is local memory, the others are CUDA-IPC mappings
  int rank = 0, world = 1;
  float* grads[DP_MAX_WORLD] = {nullptr};
  float* params[DP_MAX_WORLD] = {nullptr};
  bf16* shadow[DP_MAX_WORLD] = {nullptr};
  uint32_t* signal[DP_MAX_WORLD] = {nullptr};     // [DP_MAX_SITES][DP_MAX_WORLD] epochs per rank
  // NVLS multicast addresses (optional): one VA bound to the same offset of EVERY rank's buffer through the NVSwitch;
  // when set, the exchange kernel uses multimem.ld_reduce / multimem.st instead of per-peer loads and stores
  float* mc_grads = nullptr;
  float* mc_params = nullptr;
  bf16* mc_shadow = nullptr;
};
// diagnostics of the multicast mapping (tools/nvls_probe.py, tests): dst[i] = sum over ranks of src[i] / every rank's dst[i] = src[i]
int multimem_reduce_f32(const float* mc_src, float* dst_local, long long n, cudaStream_t st);
int multimem_bcast_f32(float* mc_dst, const float* src_local, long long n, cudaStream_t st);
int dp_barrier(const DpPeers& peers, uint32_t* epoch, uint32_t* status, int site, cudaStream_t st);
int dp_adamw_step(const DpPeers& peers, const AdamWArgs& a, long long off, cudaStream_t st);
// action_mean / action_std (optional, fp32 [A]): out = raw * std + mean in float64 (post_process, imitate_episodes.py:294)
int temporal_agg_step(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A, double k,
                      cudaStream_t st, const float* action_mean = nullptr, const float* action_std = nullptr);
// open-loop chunk replay (imitate_episodes.py:415): out[e] = chunks[e, j] with the real-robot base re-timing of :401
// (the last two action dims are read base_delay steps later), optionally de-normalised; j = t[e] % (nq - base_delay)
int chunk_select(const float* chunks, const int* t, double* out, int E, int nq, int A, int base_delay,
                 const float* action_mean, const float* action_std, cudaStream_t st);

}  // namespace act
