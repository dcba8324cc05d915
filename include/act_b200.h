/* act_b200 -- C ABI of the B200-native ACT (DETRVAE) hot path.
 *
 * Plain C: raw device pointers, sizes, an explicit cudaStream_t passed as void*.  Every function returns 0 on
 * success and a negative code on failure; act_last_error() gives the message (no exceptions cross the ABI).
 * The reference (jie0530/act-plus-plus) is pure Python/PyTorch and has no FFI; each entry point below names the
 * reference code (file:line under /root/reference) whose work it replaces.  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add.
 *
 * Layout conventions: bf16 tensors are raw uint16 storage; activations are NHWC; token tensors are [B, L, d].
 */
#ifndef ACT_B200_H_
#define ACT_B200_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* act_last_error(void);
int act_abi_version(void);

/* ---- dense contractions (tcgen05 / TMA).  Replace torch nn.Linear / bmm calls of transformer.py:194-259 and
 *      detr_vae.py:57-63 (F.linear -> cuBLAS in the reference). ---------------------------------------------- */
typedef struct {
  uint32_t flags;            /* ACT_EPI_* */
  float alpha;
  const float* bias;         /* [N] */
  const float* scale;        /* [N] */
  const float* res_f32;
  const void* res_bf16;
  const void* aux_bf16;
  float* out_f32;
  void* out_bf16;
  int64_t ldc, ldr, osz0, osz1, rsz0, rsz1;
} act_epilogue;

enum {
  ACT_EPI_BIAS = 1, ACT_EPI_SCALE = 2, ACT_EPI_RES_F32 = 4, ACT_EPI_RES_BF16 = 8, ACT_EPI_RELU = 16,
  ACT_EPI_MASK_POS = 32, ACT_EPI_OUT_F32 = 64, ACT_EPI_OUT_BF16 = 128, ACT_EPI_ATOMIC = 256
};

/* D[z][M,N] = epilogue(alpha * A[z][M,K] * B[z][N,K]^T), bf16 operands, fp32 accumulation in TMEM.
 * a_mn / b_mn = 0: operand stored K-major ((row,k) at ptr + row*ld + k); 1: MN-major ((row,k) at ptr + k*ld + row).
 * Batch z = z0 + nz0*z1 adds z0*sz0 + z1*sz1 elements to each operand. */
int act_tc_gemm(int M, int N, int K, int nz0, int nz1,
                const void* A, int a_mn, int64_t lda, int64_t a_sz0, int64_t a_sz1,
                const void* B, int b_mn, int64_t ldb, int64_t b_sz0, int64_t b_sz1,
                const act_epilogue* epi, int ksplit, int bn_hint, int stages_hint, void* stream);


/* ---- convolutions over NHWC bf16 (implicit GEMM on tcgen05, TMA tap boxes).  Replace torchvision resnet18's
 *      nn.Conv2d + FrozenBatchNorm2d (backbone.py:47-57,121-124) and their autograd backward (cuDNN in the
 *      reference).  Kernels are 1x1 or 3x3, stride 1 or 2.
 *      wpk  : bf16 [Cout][(kh*KW+kw)*Cin + ci]           (forward operand)
 *      wpkt : bf16 [Cin][(kh*KW+kw)*Cout + co] * scale   (data-gradient operand)                         ---- */
int act_cast_bf16(const float* src, void* dst, int64_t n, void* stream);
int act_pack_conv_weights(const float* w, const float* scale, void* wpk, void* wpkt, int Cout, int Cin, int KK,
                          void* stream);
/* y[N,Ho,Wo,Cout] = epilogue(conv(x[N,H,W,Cin], w)); epilogue geometry (ldc etc.) is implied by the shape. */
int act_conv_fwd(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x, const void* wpk,
                 const act_epilogue* epi, void* stream);
/* dx[N,H,W,Cin] = epilogue(conv_transpose(dy[N,Ho,Wo,Cout], w)); for stride 2 `res_compact` ([N,H/2,W/2,Cin] bf16,
 * may be NULL) is added on the even/even pixels and epi's own residual is ignored. */
int act_conv_dgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* dy,
                   const void* wpkt, const act_epilogue* epi, const void* res_compact, void* stream);
/* dw[Cout][Cin][KH][KW] (fp32) += alpha * scale[co] * sum_pixels dy * x  (atomic accumulation). */
int act_conv_wgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x, const void* dy,
                   float* dw, float alpha, const float* scale, void* stream);
/* Allocation-free variant of act_conv_wgrad.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments and result; the [KH*KW][Cout][Cin] fp32 staging tile lives in a static device array of 589824 floats
 * (2.25 MiB per GPU that loads the library, held until exit), so KH*KW*Cin*Cout <= 9*256*256 (every resnet18 conv up to
 * layer3); larger shapes return nonzero with a message in act_last_error().  The call performs no cudaMalloc / cudaFree.
 * All calls share the array (each re-zeroes the part it uses on `stream`): calls that overlap on the GPU race. */
#define ACT_CONV_WGRAD_NOALLOC_MAX_FLOATS 589824
int act_conv_wgrad_noalloc(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x,
                           const void* dy, float* dw, float alpha, const float* scale, void* stream);

/* ---- fused bandwidth-bound kernels (warp-shuffle reductions, 128-bit accesses) -------------------------------- */
/* nn.LayerNorm(d) of transformer.py:200-201,261-263,35: y = LN(t)*gamma+beta (eps 1e-5), d in {256,512}.  Optional
 * outputs: y fp32, bf16(y), bf16(y + pos[row % L]) (with_pos_embed, transformer.py:205-206), mean, rstd. */
int act_ln_fwd(const float* t, const float* gamma, const float* beta, const float* pos, int L, float* y, void* y_bf16,
               void* ypos_bf16, float* mean, float* rstd, int M, int D, void* stream);
/* backward of the above for dy (+ optional second addend dy2); dgamma / dbeta are accumulated (atomics). */
int act_ln_bwd(const float* dy, const float* dy2, const float* t, const float* mean, const float* rstd,
               const float* gamma, float* dt, void* dt_bf16, float* dgamma, float* dbeta, int M, int D, void* stream);
int act_add_pos_cast(const float* x, const float* pos, int L, void* x_bf16, void* xpos_bf16, int M, int D, void* stream);
/* softmax over keys of nn.MultiheadAttention (transformer.py:194,254-255): S fp32 [R, Lp] -> P bf16 [R, Lp];
 * key_padding_mask uint8 [B, L] (may be NULL), batch of row r is r / rows_per_batch; columns >= L are written 0. */
int act_softmax_fwd(const float* S, const uint8_t* key_padding_mask, void* P, int R, int L, int Lp, int rows_per_batch,
                    void* stream);
int act_softmax_bwd(const void* P, const float* dP, void* dS, int R, int L, int Lp, float scale, void* stream);
/* Attention scores fused with softmax / softmax backward (head_dim 64, Lk <= 320; nn.MultiheadAttention's
 * softmax(q k^T / 8 + key_padding_mask) `transformer.py:194,254-255` and its autograd):
 *   bwd = 0: out[b,h,i,:] = softmax_j(scale * sum_d A[b,h,i,d] B[b,h,j,d] + mask[b,j])      (A = Q, B = K)
 *   bwd = 1: out = scale * P o (dP - rowsum(dP o P)),  dP[b,h,i,j] = sum_d A[b,h,i,d] B[b,h,j,d]   (A = dO, B = V)
 * A / B are bf16 row views (row stride ld*, head stride *_sz0, batch stride *_sz1, in elements); P / out are bf16
 * [batch, heads, Lq, Lkp], columns [Lk, Lkp) are written 0; mask is uint8 [batch, Lk] or NULL. */
int act_attn_scores(int bwd, const void* A, int64_t lda, int64_t a_sz0, int64_t a_sz1, const void* B, int64_t ldb, int64_t b_sz0,
                    int64_t b_sz1, int Lq, int Lk, int Lkp, int heads, int batch, float scale, const uint8_t* key_padding_mask,
                    const void* P, void* out, void* stream);
/* The bwd = 1 form with the forward's attention output O = P V given as well (row view like dO): rowsum(dP o P) equals
 * dO_i . O_i, so rows of more than 320 keys need one pass over the keys instead of two.  O == NULL: same as above. */
int act_attn_scores_bwd(const void* dO, int64_t ld_do, int64_t do_sz0, int64_t do_sz1, const void* V, int64_t ldv, int64_t v_sz0,
                        int64_t v_sz1, const void* O, int64_t ldo, int64_t o_sz0, int64_t o_sz1, int Lq, int Lk, int Lkp, int heads,
                        int batch, float scale, const void* P, void* dS, void* stream);
int act_colsum_bf16(const void* X, int64_t ld, int M, int N, float* out_accum, void* stream);
/* fp32 mini-GEMM for the K<=32 / N<=64 projections (detr_vae.py:57-63,81-97) and their gradients:
 * C(i,j) (+)= alpha*sum_r A(i,r)B(r,j) (+bias[j]); each index maps as (idx/rpb)*bs + (idx%rpb)*s (maps = 3 int64). */
int act_sgemm(int I, int J, int R, const float* A, const int64_t* a_i, const int64_t* a_r, const float* B,
              const int64_t* b_r, const int64_t* b_j, float* C, const int64_t* c_i, const int64_t* c_j,
              const float* bias, float alpha, int accumulate, void* stream);
/* L1 + KL loss of policy.py:314-318,378-391 -> losses[3] = {l1, kl, loss}; optional gradients (scaled by gscale). */
int act_loss(const float* a_hat, const float* actions, int64_t actions_batch_stride, const uint8_t* is_pad,
             int64_t is_pad_batch_stride, const float* mu, const float* logvar, int B, int nq, int A, int Z,
             float kl_weight, float gscale, float* losses, float* d_a_hat, float* d_mu, float* d_logvar, void* stream);
/* reparametrize of detr_vae.py:19-22 with injected noise: latent_info [B,2Z] -> mu, logvar, z = mu + exp(lv/2)*eps */
int act_reparam_fwd(const float* latent_info, const float* eps, float* mu, float* logvar, float* z, int B, int Z,
                    void* stream);
int act_reparam_bwd(const float* dz, const float* eps, const float* logvar, const float* d_mu_kl,
                    const float* d_logvar_kl, float* d_latent_info, int B, int Z, void* stream);
/* resnet stem: max-pool 3x3/s2/p1 on NHWC bf16 (+arg-max bytes) and its ReLU-fused backward (torchvision resnet18
 * `maxpool`, backbone.py:121-124).  The stem conv itself is an implicit GEMM inside the engine (no im2col).
 * argmax holds the winning tap 0..8 (first maximum in kh,kw order, as torch), or 9 when the window maximum is not positive
 * (nothing passes the fused ReLU); act_maxpool_bwd therefore accepts x == NULL and then takes the ReLU mask from the record
 * alone (the engine's path: 79 MB less to read at batch 8); with x given it also applies (x > 0) explicitly. */
int act_maxpool_fwd(const void* x, void* y, uint8_t* argmax, int N, int H, int W, int C, void* stream);
int act_maxpool_bwd(const void* dy, const uint8_t* argmax, const void* x, void* dx, int N, int H, int W, int C,
                    void* stream);
/* torch.optim.AdamW single step over one flat segment (detr/main.py:102-110); mode 1 = decay only (zero gradient). */
int act_adamw_step(float* p, const float* g, float* m, float* v, void* shadow_bf16, int64_t n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int step, int mode, void* stream);
/* temporal ensemble of imitate_episodes.py:339,403-411 for E environments at once: ring fp32 [E,nq,nq,A],
 * chunks fp32 [E,nq,A] (the policy output at step t[e]), out float64 [E,A]. */
int act_temporal_agg_step(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A, double k,
                          void* stream);

/* Same with the de-normalisation of imitate_episodes.py:294 fused (post_process: action = raw * std + mean, float64 raw
 * with float32 statistics): action_mean / action_std fp32 [A] device pointers (both or neither). */
int act_temporal_agg_step_ex(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A, double k,
                             const float* action_mean, const float* action_std, void* stream);
/* Open-loop chunk replay (imitate_episodes.py:415 `raw_action = all_actions[:, t % query_frequency]`) with the real-robot
 * base re-timing of :311-313, 401 (query_frequency = nq - base_delay; the last two action dims are read base_delay steps
 * later: cat([a[:, :-D, :-2], a[:, D:, -2:]])) and the optional de-normalisation: out float64 [E, A]. */
int act_chunk_select(const float* chunks, const int* t, double* out, int E, int nq, int A, int base_delay,
                     const float* action_mean, const float* action_std, void* stream);

/* ---- the policy engine: DETRVAE forward / backward / AdamW as one stream-ordered object -----------------------
 * Replaces detr/main.py:92-112 (build_ACT_model_and_optimizer), ACTPolicy.__call__ (policy.py:264-332),
 * loss.backward() / optimizer.step() of imitate_episodes.py:603-607.  Parameters, gradients and Adam moments are
 * caller-owned flat fp32 device buffers (torch tensors on the Python side); the tensor table tells where each
 * reference state_dict entry lives inside them.  Workspace is owned by the handle. */
typedef struct act_engine act_engine;
typedef struct {
  int batch_max, num_cameras, img_h, img_w;
  int hidden_dim, nheads, dim_feedforward, enc_layers, dec_layers;
  int num_queries, state_dim, action_dim, latent_dim;
  float kl_weight, lr, lr_backbone, weight_decay, beta1, beta2, adam_eps;
  int no_encoder;         /* 1: DETRVAE built with encoder=None (detr_vae.py:109-112, 373-377): no `encoder.*` tensors,
                            latent = 0 in every call, kl = 0 (policy.py:306-307) */
  int train;              /* 0: inference-only workspace */
} act_engine_config;

int act_engine_create(const act_engine_config* cfg, act_engine** out);
void act_engine_destroy(act_engine* e);
int act_engine_num_tensors(const act_engine* e);
int64_t act_engine_total_floats(const act_engine* e);
int64_t act_engine_workspace_bytes(const act_engine* e);
/* kind: 0 param, 1 param with identically-zero grad (dead decoder layers), 2 param without grad (is_pad_head),
 * 3 backbone param (lr_backbone), 4 buffer.  region: 0 main, 1 cvae, 2 dead, 3 nograd, 4 backbone, 5 buffer. */
int act_engine_tensor_info(const act_engine* e, int i, char* name, int name_cap, int64_t shape[4], int* ndim,
                           int64_t* offset, int64_t* numel, int* kind, int* region);
int act_engine_region(const act_engine* e, int region, int64_t* offset, int64_t* count);
int act_engine_bind(act_engine* e, float* params, float* grads, float* exp_avg, float* exp_avg_sq);
/* rebuild the bf16 / packed operand shadows from the fp32 parameters (after load_state_dict or any external write) */
int act_engine_refresh(act_engine* e, void* stream);
/* qpos [B,S] f32, image [B,K,3,H,W] f32 in [0,1]; training/validation call when actions != NULL
 * (actions [B,>=nq,A] with batch stride actions_bs, is_pad uint8 [B,>=nq] with batch stride is_pad_bs, eps [B,Z]).
 * Outputs (each may be NULL): a_hat [B,nq,A], is_pad_hat [B,nq], mu/logvar [B,Z], losses[3] = {l1, kl, loss}. */
int act_engine_forward(act_engine* e, int B, const float* qpos, const float* image, const float* actions,
                       int64_t actions_bs, const uint8_t* is_pad, int64_t is_pad_bs, const float* eps, float* a_hat,
                       float* is_pad_hat, float* mu, float* logvar, float* losses, void* stream);
/* gradients of loss_grad * loss into the bound gradient buffer.  stage 0 = everything; 1 = heads + transformer,
 * 2 = CVAE encoder, 3 = backbones, 4 = 2 and 3 as two concurrent branches (an internal side stream forked from and
 * joined into `stream`).  Run 1 then 4 (or 1,2,3) to overlap the gradient all-reduce of finished regions. */
int act_engine_backward(act_engine* e, int stage, float loss_grad, void* stream);
int act_engine_zero_grad(act_engine* e, void* stream);
/* same, but the clear runs beside the following forward on an internal stream; the next act_engine_backward /
 * act_engine_backward_and_step / act_engine_adamw_step on this engine waits for it. */
int act_engine_zero_grad_async(act_engine* e, void* stream);
/* loss.backward() + optimizer.step() (imitate_episodes.py:603-607) as one call: act_engine_backward(e, 0, ...) followed
 * by act_engine_adamw_step(e, step, ...), with each parameter region's AdamW launched as soon as that region's
 * gradients are final so it overlaps the remaining backward.  Single-process only (no gradient all-reduce in between). */
int act_engine_backward_and_step(act_engine* e, int step, float loss_grad, void* stream);
/* Data-parallel form of the call above (new work: the reference trains on one GPU).  While the step is
 * being enqueued the engine calls grads_ready(user, offset, count, s) once for every contiguous range of the flat
 * fp32 gradient buffer (act_engine_grads) as soon as the kernels producing it are enqueued; `s` is the CUDA stream on
 * which the range is final.  The callback must enqueue the cross-rank average of grads[offset, offset+count) ordered
 * after `s` and make `s` wait for it (NCCL all-reduce + event wait), then return 0; the engine enqueues the range's
 * AdamW on `s` right behind it.  The ranges cover the main, CVAE-encoder and backbone regions exactly once per step,
 * in the same order on every rank, so each collective overlaps the backward work that is still running. */
typedef int (*act_grads_ready_fn)(void* user, long long offset, long long count, void* stream);
int act_engine_backward_and_step_dp(act_engine* e, int step, float loss_grad, void* stream,
                                    act_grads_ready_fn grads_ready, void* user);
/* Data-parallel step over NVLink peer memory (no NCCL on the data path).  Set-up, once per process:
 *   act_engine_dp_export   writes ACT_DP_BLOB_BYTES of CUDA-IPC handles (flat gradients, flat parameters, bf16 shadow,
 *                          barrier flags) of this rank; exchange the blobs between ranks by any means (the Python host
 *                          uses torch.distributed.all_gather_object);
 *   act_engine_dp_connect  maps every peer's buffers (blobs = world x ACT_DP_BLOB_BYTES, rank order);
 *   act_engine_dp_set_peers is the same with raw device pointers that are already valid on this device (engines of one
 *                          process; entry `rank` must be this engine's own buffers as returned by act_engine_dp_local).
 * act_engine_backward_and_step_p2p then runs the step of act_engine_backward_and_step with every "all-reduce the range,
 * AdamW the range" pair replaced by a flag barrier in peer memory and ONE kernel per rank that reduces its 1/world slice
 * of the range from all ranks' gradient buffers, applies AdamW with its private moment slice and stores the new fp32
 * parameters + bf16 shadow into every rank's buffers (reduce-scatter + optimiser + all-gather in one pass).  Moments
 * are therefore sharded: rank r holds the current exp_avg / exp_avg_sq only for its slices (act_engine_dp_slices), and
 * the local gradient buffer keeps this rank's un-averaged gradients.  Every rank must issue the same sequence of
 * calls.  act_engine_dp_status returns 1 if a barrier timed out (a peer never arrived); it synchronises the device. */
#define ACT_DP_BLOB_BYTES 296
int act_engine_dp_local(act_engine* e, void** grads, void** params, void** shadow, void** signal);
int act_engine_dp_export(act_engine* e, void* blob, int capacity);
int act_engine_dp_connect(act_engine* e, int rank, int world, const void* blobs, int blob_bytes);
int act_engine_dp_set_peers(act_engine* e, int rank, int world, void* const* grads, void* const* params,
                            void* const* shadow, void* const* signal);
/* NVLS (NVSwitch multicast) form of the same exchange, no reference counterpart: when the four exchange buffers live in
 * symmetric memory that also has a multicast mapping (the Python host takes it from torch.distributed._symmetric_memory:
 * plumbing only), the kernel of act_engine_backward_and_step_p2p loads the switch-reduced gradient with ONE
 * multimem.ld_reduce and stores parameters + shadow to all ranks with ONE multimem.st each, so every rank moves
 * 4 + 6 B per owned parameter over NVLink instead of (world-1) * (4 + 6) B.
 *   act_engine_dp_use_buffers  (before act_engine_dp_set_peers) moves the bf16 shadow into caller-owned memory and adopts a
 *                              caller-owned flag block of act_engine_dp_signal_bytes() bytes;
 *   act_engine_dp_set_multicast (after act_engine_dp_set_peers) gives the multicast addresses of the gradient / parameter /
 *                              shadow buffers (all three or, to switch back to per-peer accesses, three NULLs).
 * act_multimem_reduce_f32 / act_multimem_bcast_f32 are diagnostics of such a mapping (dst = sum over ranks of src / every
 * rank's dst = src; n a multiple of 4, 16-byte aligned). */
int act_engine_dp_use_buffers(act_engine* e, void* shadow, void* signal);
int64_t act_engine_dp_signal_bytes(void);
int act_engine_dp_set_multicast(act_engine* e, void* mc_grads, void* mc_params, void* mc_shadow);
int act_multimem_reduce_f32(const float* mc_src, float* dst_local, int64_t n, void* stream);
int act_multimem_bcast_f32(float* mc_dst, const float* src_local, int64_t n, void* stream);
int act_engine_backward_and_step_p2p(act_engine* e, int step, float loss_grad, void* stream);
int act_engine_dp_status(act_engine* e);
/* Marks peer `peer_rank` as having arrived at every barrier up to `epoch` (writes this engine's own flag block).  For
 * single-process tests and dry runs in which a peer is a plain set of buffers rather than a running engine. */
int act_engine_dp_preset_signal(act_engine* e, int peer_rank, unsigned int epoch);
/* The [offset, offset+count) ranges (in floats of the flat buffers) that the last p2p step exchanged, in order; returns
 * the number of ranges (at most `capacity` are written).  Rank r owns floats [4*lo, 4*hi) of a range with
 * n4 = count/4, lo = r*(n4/world) + min(r, n4%world), hi = lo + n4/world + (r < n4%world); rank 0 also owns the
 * count%4 tail. */
int act_engine_dp_slices(act_engine* e, long long* offsets, long long* counts, int capacity);
/* AdamW of one parameter region (index as in act_engine_region; the backbone region also re-packs the conv operands).
 * blocks_per_sm > 0 limits the launch to that many 256-thread blocks per SM so it can run beside tensor-core kernels.
 * The data-parallel host loop uses it to update a region as soon as its gradient all-reduce has finished. */
int act_engine_adamw_region(act_engine* e, int step, int region, int blocks_per_sm, void* stream);
/* on = 0: every launch of this engine goes to the caller's stream in program order (no internal branches); used to time
 * kernels one at a time.  Call only between steps (never between zero_grad_async / forward / backward of one step). */
int act_engine_set_concurrent(act_engine* e, int on);
/* Layout of the `image` argument of the following act_engine_forward calls: 0 = float32 [B, K, 3, H, W] in [0, 1] (the
 * reference call signature, policy.py:264); 1 = uint8 [B, K, H, W, 3], the dataset's native frames (utils.py:133-158
 * divides by 255 and permutes on the host) -- the /255, the ImageNet normalisation and the layout change are fused into
 * the stem's image pack, and the host->device copy shrinks 4x. */
int act_engine_set_image_format(act_engine* e, int fmt);
int act_engine_adamw_step(act_engine* e, int step, void* stream);
int act_engine_set_lr(act_engine* e, float lr, float lr_backbone);
/* Train-mode dropout of the transformer blocks (transformer.py:194,197,202-203,219-222,254-266,284-293; p = 0.1 at
 * detr/main.py:45; nn.MultiheadAttention's attention-weight dropout included).  p = 0 switches it off (`.eval()`).
 * Masks are a counter-based function Philox4x32-10(seed; element / 8, site, step) regenerated wherever needed -- nothing
 * is stored; `step` is a device counter bumped by every training forward (CUDA-graph replays draw fresh masks).
 * Data-parallel ranks pass different seeds.  Applies to act_engine_forward calls WITH actions (and their backward). */
/* Rollout pre-processing on the device (imitate_episodes.py:290, 370 `pre_process`): inference calls (actions == NULL)
 * z-score the raw qpos inside the proprio projection, (qpos - mean) / std in fp32 like the reference; NULL, NULL = off. */
int act_engine_set_qpos_norm(act_engine* e, const float* mean, const float* std, void* stream);
/* host-side switch over the stored statistics (no device work): a rollout object turns it on around its own queries */
int act_engine_enable_qpos_norm(act_engine* e, int on);
int act_engine_set_dropout(act_engine* e, float p, uint64_t seed);
int act_engine_dropout_state(act_engine* e, float* p, uint64_t* seed, uint32_t* step);   /* synchronises */
/* The keep mask (uint8, 1 = kept) of elements [0, n) of dropout site `site` at `step`: the same function the kernels
 * evaluate.  Site ids: 0x100 + 8 l + k main encoder layer l, 0x200 + 8 l + k CVAE encoder, 0x300 + k decoder layer 0;
 * k = 0 attention weights [B, heads, Lq, Lkp], 1 attention out-proj [B L, d], 2 FFN hidden [B L, ff], 3 FFN out [B L, d],
 * 4 / 5 decoder cross-attention weights / out-proj. */
int act_dropout_mask(uint64_t seed, uint32_t site, uint32_t step, float p, int64_t n, uint8_t* out, void* stream);
/* loss.backward() hands autograd's incoming gradient over ON THE DEVICE: gvec3 = d/d{l1, kl, loss} (three floats; the
 * dict of policy.py:316-318 returns all three scalars and any of them may be the root).  The following
 * act_engine_backward* calls read it inside the loss kernel instead of the scalar loss_grad argument -- no host
 * synchronisation in the step.  NULL restores the scalar argument. */
int act_engine_set_loss_grad(act_engine* e, const float* gvec3);
/* `stream` waits for every internal lane of the engine (an act_engine_zero_grad_async clear, trailing weight-gradient
 * launches): closes a CUDA-graph capture that ends between the forward and the backward of a step. */
int act_engine_join_lanes(act_engine* e, void* stream);
int act_engine_launches(act_engine* e, int reset);
/* Parity diagnostics: copy of an internal activation of the LAST forward.  name: "src" (encoder input, fp32
 * [B, L, d]), "memory" (encoder output, fp32 [B, L, d]), "hs" (decoder output after the final norm, fp32 [B, nq, d]),
 * "feat<k>" (camera k's layer4 output, bf16 NHWC [B, Hf, Wf, 512]).  dtype: 0 = fp32, 1 = bf16.  dst = NULL only
 * queries numel / dtype; otherwise numel elements are copied to the device buffer dst (capacity in bytes) on `stream`. */
int act_engine_debug_buffer(act_engine* e, const char* name, void* dst, int64_t capacity, int64_t* numel, int* dtype,
                            void* stream);

/* ---- measurement hook: CUDA-event timing of every tcgen05 kernel launched between begin and end on its own stream;
 *      returns summed device time, summed algorithmic FLOPs (2*M*N*K per launch), launch count, longest launch. */
/* Differential timing: skip every launch of the selected tensor-core kernel classes (bit 0 GEMM, 1 tap-table conv,
 * 2 halo-resident 3x3 conv, 3 conv weight gradient, 4 fused attention scores).  Timing a serial CUDA-graph replay of the
 * step with and without a class gives that class's device time without per-launch events.  Outputs are garbage while a
 * mask is set.  act_tc_skipped_flops: algorithmic FLOPs of the launches skipped since the mask was last set. */
int act_tc_set_skip_mask(uint32_t mask);
double act_tc_skipped_flops(void);
/* Diagnostics: while buf != NULL, CTA 0 of every tc_kernel launch writes eight %globaltimer stamps (entry, set-up done,
 * dependency wait over, first operands landed, last MMA issued, accumulator ready, epilogue done, exit) to
 * buf[1 + 8 * n ..] with n = atomic counter buf[0] (uint64 device memory, zeroed by the caller). */
int act_tc_debug_timestamps(void* buf);
int act_tc_profile_begin(void);
int act_tc_profile_set_dump(const char* csv_path);   /* optional per-launch CSV written by act_tc_profile_end */
int act_tc_profile_end(double* total_ms, double* total_flops, int* launches, double* max_ms);

#ifdef __cplusplus
}
#endif
#endif /* ACT_B200_H_ */
