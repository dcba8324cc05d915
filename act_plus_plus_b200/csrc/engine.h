// ACT DETRVAE engine: owns the parameter layout, bf16 operand shadows, activation workspace and the kernel
// sequences of the forward (train / inference), backward and optimizer steps.  One engine per device per process;
// all calls are stream-ordered and not thread-safe (the reference is single-threaded too, imitate_episodes.py:561-612).
#pragma once
#include <map>
#include <string>
#include <vector>
#include "conv_ops.h"
#include "kernels.h"

namespace act {

struct EngineConfig {
  int batch_max = 8;
  int num_cameras = 1;
  int img_h = 480, img_w = 640;
  int hidden_dim = 512, nheads = 8, dim_ff = 3200, enc_layers = 4, dec_layers = 7;
  int num_queries = 100, state_dim = 14, action_dim = 16, latent_dim = 32;
  float kl_weight = 10.f;
  float lr = 1e-5f, lr_backbone = 1e-5f, weight_decay = 1e-4f, beta1 = 0.9f, beta2 = 0.999f, adam_eps = 1e-8f;
  int no_encoder = 0;         // 1: encoder=None variant (detr_vae.py:109-112): latent 0, kl 0, no `encoder.*` tensors
  int train = 1;              // 0: inference-only engine (no backward workspace)
};

enum TensorKind { T_PARAM = 0, T_PARAM_DEAD = 1, T_PARAM_NOGRAD = 2, T_PARAM_BACKBONE = 3, T_BUFFER = 4 };
enum Region { R_MAIN = 0, R_CVAE = 1, R_DEAD = 2, R_NOGRAD = 3, R_BACKBONE = 4, R_BUFFER = 5, R_COUNT = 6 };

struct TensorEntry {
  std::string name;
  int ndim = 0;
  long long shape[4] = {1, 1, 1, 1};
  long long numel = 0, offset = 0;   // offset in floats into the flat buffers
  int kind = T_PARAM;
  int region = R_MAIN;
  int order = 0;                     // position in the reference's state_dict order
};

struct LinW { const bf16* w = nullptr; const float* b = nullptr; const float* w32 = nullptr; float* gw = nullptr; float* gb = nullptr; int N = 0, K = 0; };
struct LnW { const float* g = nullptr; const float* b = nullptr; float* gg = nullptr; float* gb = nullptr; };
struct MhaW { LinW qk, q, k, v, o; };
struct EncW { MhaW sa; LinW l1, l2; LnW n1, n2; };
struct DecW { MhaW sa, ca; LinW l1, l2; LnW n1, n2, n3; };
struct ConvW {
  int Cin = 0, Cout = 0, KH = 0, stride = 1, pad = 0;
  const float* w32 = nullptr; float* gw = nullptr;
  const float* bn[4] = {nullptr, nullptr, nullptr, nullptr};   // weight, bias, running_mean, running_var
  bf16* wpk = nullptr; bf16* wpkt = nullptr;                   // packed shadows (engine-owned)
  float* scale = nullptr; float* shift = nullptr;              // folded FrozenBN (engine-owned)
  // 64 -> 64 channel 3x3/s1 convs only (ResNet layer1): paired operands [128][9*128] (conv_ops.cu::pair_launch) and the
  // folded FrozenBN repeated to 128 entries
  bf16* wpk2 = nullptr; bf16* wpkt2 = nullptr; float* scale2 = nullptr; float* shift2 = nullptr;
};

struct BlockW { ConvW c1, c2, ds; bool has_ds = false; };
struct CamW { ConvW stem; BlockW blk[8]; };

struct EncActs {       // saved activations of one post-norm encoder layer over M = B*L rows
  bf16 *xb = nullptr, *xpb = nullptr, *qk = nullptr, *v = nullptr, *P = nullptr, *o = nullptr, *x1b = nullptr, *hh = nullptr;
  bf16* P2 = nullptr;    // dropout(P), the operand of P.V in train mode (P itself feeds the softmax backward)
  float *t1 = nullptr, *mean1 = nullptr, *rstd1 = nullptr, *x1 = nullptr, *t2 = nullptr, *mean2 = nullptr, *rstd2 = nullptr, *x2 = nullptr;
};
struct DecActs {
  bf16 *xb = nullptr, *xpb = nullptr, *qk = nullptr, *v = nullptr, *P = nullptr, *o = nullptr;     // self-attention
  bf16 *P2 = nullptr, *cP2 = nullptr;                                                               // dropped attention weights
  float *t1 = nullptr, *mean1 = nullptr, *rstd1 = nullptr, *y1 = nullptr; bf16* y1pb = nullptr;
  bf16 *q = nullptr, *k = nullptr, *cv = nullptr, *cP = nullptr, *co = nullptr;                     // cross-attention
  float *t2 = nullptr, *mean2 = nullptr, *rstd2 = nullptr, *y2 = nullptr; bf16* y2b = nullptr;
  bf16* hh = nullptr;
  float *t3 = nullptr, *mean3 = nullptr, *rstd3 = nullptr, *y3 = nullptr;
};
struct BlockActs { bf16 *a = nullptr, *y = nullptr, *ds = nullptr; };   // conv1 relu out, block out, downsample out
struct CamActs {
  bf16 *xp = nullptr;                                                   // normalised NHWC4 image, padded rows [B][H][Wp][4]
  bf16 *y1 = nullptr, *p1 = nullptr; uint8_t* pidx = nullptr;           // stem
  BlockActs blk[8];
};

typedef int (*GradsReadyFn)(void* user, long long offset, long long count, void* stream);

class Engine {
 public:
  explicit Engine(const EngineConfig& cfg);
  ~Engine();
  int init();                                   // builds the tensor table, allocates workspace
  // parameter plumbing
  const std::vector<TensorEntry>& tensors() const { return tensors_; }
  long long total_floats() const { return total_floats_; }
  void region(int r, long long* off, long long* n) const { *off = region_off_[r]; *n = region_n_[r]; }
  int bind(float* params, float* grads, float* exp_avg, float* exp_avg_sq);
  int refresh_shadows(cudaStream_t st);         // fp32 params -> bf16 mirrors / packed conv weights / BN folds / pos rows
  // steps
  int forward(int B, const float* qpos, const float* image, const float* actions, long long actions_bs,
              const uint8_t* is_pad, long long is_pad_bs, const float* eps, float* a_hat, float* is_pad_hat, float* mu,
              float* logvar, float* losses, cudaStream_t st);
  int backward(int stage, float loss_grad, cudaStream_t st);      // stage: 0 = all, 1 = transformer, 2 = cvae, 3 = backbone, 4 = cvae || backbone
  int zero_grad(cudaStream_t st, bool async = false);
  int adamw(int step, cudaStream_t st, bool begin = true);
  int adamw_region(int r, int step, cudaStream_t st, int blocks_per_sm = 0);
  int adamw_range(long long off, long long n, float lr, bool shadow, int step, cudaStream_t st, int blocks_per_sm);
  GradsReadyFn grads_cb_ = nullptr;
  void* grads_cb_user_ = nullptr;
  int grads_ready(long long off, long long n, cudaStream_t st);
  int backbone_bwd_and_step(cudaStream_t st);
  int opt_step_ = 0, dp_bb_slot_ = 0;
  bool dp_bb_pending_ = false;
  cudaStream_t dp_main_st_ = nullptr;
  int dp_inline_from_ = 0;                       // CVAE layers >= this index are updated inside cvae_bwd, the rest at its end
  bool inline_opt_on_ = false;
  int inline_opt_step_ = 0;                      // > 0 while backward_and_step runs: cvae_bwd updates each layer as soon as its gradients are final
  long long cvae_layer_off_[16] = {0}, cvae_layer_n_[16] = {0};
  // per camera: the layer4 conv weights (75 % of a ResNet-18's parameters) are a contiguous tail of the camera's span in
  // R_BACKBONE.  Their gradients are final a quarter into the backbone backward (it runs layer4 -> stem), so inside
  // backward_and_step they are folded, exchanged and updated early; only the small remainder trails the branch.
  long long bb_l4_off_[8] = {0}, bb_l4_n_[8] = {0}, bb_rest_off_[8] = {0}, bb_rest_n_[8] = {0};
  bool bb_early_ = false;                        // set while backbone_bwd_and_step runs
  bool bb_split_ = false;                        // single-GPU step: AdamW + re-pack of everything but conv1 right after the fold
  // backward(0) + adamw() with the optimiser of each parameter region trailing that region's last weight gradient, so
  // the HBM-bound AdamW of the transformer / CVAE regions overlaps the tensor-bound backbone backward
  // grads_ready: data-parallel hook.  Called on the host while the step is enqueued, once per contiguous range of the
  // flat gradient buffer, with the stream on which that range is final; it must enqueue the cross-rank average of the
  // range and order `stream` after it.  The range's AdamW follows on the same stream.
  int backward_and_step(int step, float loss_grad, cudaStream_t st, GradsReadyFn cb = nullptr, void* cb_user = nullptr,
                        bool p2p = false);
  // ---- data-parallel step over NVLink peer memory (dp_p2p.cu): every rank maps the other ranks' flat gradient /
  // parameter / bf16-shadow buffers and a block of barrier flags; backward_and_step(p2p = true) then replaces each
  // "all-reduce the range, AdamW the range" pair by a flag barrier and ONE kernel that reduces this rank's slice of the
  // range from all ranks, updates it and stores the new parameters into every rank.  All of it runs on one dedicated
  // stream in a fixed order (same on every rank), so the spinning barrier kernels cannot wait on each other crosswise.
  static constexpr int DP_BLOB_BYTES = 4 * 64 + 4 * 8 + 8;
  int dp_local(void** grads, void** params, void** shadow, void** signal);      // this rank's four buffers (allocates the flags)
  int dp_export(void* blob, int cap);                                             // CUDA-IPC handles of the four buffers
  int dp_connect(int rank, int world, const void* blobs, int blob_bytes);         // map every peer's buffers
  int dp_set_peers(int rank, int world, void* const* grads, void* const* params, void* const* shadow, void* const* signal);
  int dp_preset_signal(int peer_rank, uint32_t epoch);
  // symmetric-memory set-up (before dp_set_peers): move the bf16 shadow into caller-owned memory (contents copied) and use a
  // caller-owned, zeroed flag block, so that all four exchange buffers can come from one symmetric allocator
  int dp_use_buffers(void* shadow, void* signal);
  int dp_set_multicast(void* mc_grads, void* mc_params, void* mc_shadow);      // after dp_set_peers; nullptrs switch it off
  static constexpr long long DP_SIGNAL_BYTES = 4LL * DP_MAX_SITES * DP_MAX_WORLD;
  int dp_status();                                                                // 0 ok, 1 a barrier timed out (synchronises)
  int dp_world() const { return peers_.world; }
  int dp_slices(long long* off, long long* n, int cap) const {
    const int k = static_cast<int>(dp_ranges_.size());
    for (int i = 0; i < k && i < cap; ++i) { off[i] = dp_ranges_[i].first; n[i] = dp_ranges_[i].second; }
    return k;
  }
  const EngineConfig& cfg() const { return cfg_; }
  long long workspace_bytes() const { return ws_bytes_; }
  int launches() const { return launches_; }
  int debug_buffer(const char* name, void** ptr, long long* numel, int* dtype);
  int* opt_step_dev_ = nullptr;                 // device copy of the optimiser step (AdamW bias corrections read it)
  // step > 0: the device counter is set to it; step == 0: the counter is incremented (CUDA-graph replays advance)
  int begin_opt_step(int step, cudaStream_t st);
  // Train-mode dropout (transformer.py:194,197,202-203,219-222,254-266,284-293; detr/main.py:45).  p = 0 turns it off
  // (the reference's .eval()).  Masks are Philox functions of (seed, site, step, element) regenerated in the backward;
  // the step counter lives on the device and is bumped by every training forward (CUDA-graph replays draw new masks).
  // Rollout pre-processing on the device (imitate_episodes.py:290, 370): inference calls z-score the raw qpos inside the
  // proprio projection's operand load; mean = std = NULL switches it off.  Training calls are never touched (the dataset
  // normalises, utils.py:160-166).  The vectors are copied.
  int set_qpos_norm(const float* mean_dev, const float* std_dev, cudaStream_t st);
  float *qnorm_mean_ = nullptr, *qnorm_std_ = nullptr;
  bool qnorm_on_ = false, qnorm_set_ = false;
  int set_dropout(float p, unsigned long long seed);
  int dropout_state(float* p, unsigned long long* seed, uint32_t* step);   // synchronises
  enum : uint32_t { SITE_ENC = 0x100, SITE_CVAE = 0x200, SITE_DEC = 0x300 };   // + 8 * layer + {0 attn, 1 out-proj, 2 ffn hidden, 3 ffn out, 4 cross attn, 5 cross out-proj}
  DropDesc drop_site(uint32_t site) const;
  void reset_launches() { launches_ = 0; }
  // 0: image is float32 [B, K, 3, H, W] in [0, 1] (policy.py:264 signature); 1: uint8 [B, K, H, W, 3] (dataset frames)
  int set_image_format(int fmt) { if (fmt != 0 && fmt != 1) { set_error("image format must be 0 (f32 CHW) or 1 (u8 HWC)"); return -1; } image_fmt_ = fmt; return 0; }
  int image_fmt_ = 0;
  int set_concurrent(bool on) { if (ensure_streams()) return -1; concurrent_ = on; return 0; }
  void set_lr(float lr, float lr_backbone) { cfg_.lr = lr; cfg_.lr_backbone = lr_backbone; }
  // device pointer to the incoming gradient of {l1, kl, loss} (autograd hands it over on the device); null = the scalar
  // loss_grad argument of backward() applies to `loss` alone.  Sticky until changed.
  void set_loss_grad(const float* gvec3) { loss_gvec_ = gvec3; }
  const float* loss_gvec_ = nullptr;
  int join_lanes(cudaStream_t st) { return wg_join(st); }     // st waits for every internal lane (e.g. an async zero_grad)

 private:
  // ---- helpers (engine_fwd.cu / engine_bwd.cu)
  int linear(const bf16* X, int M, const LinW& w, uint32_t flags, const float* res_f32, float* out_f32, bf16* out_bf16,
             const bf16* aux, cudaStream_t st, bool splitk = false, uint32_t drop = 0);
  int linear_dx(const bf16* dY, int M, const LinW& w, uint32_t flags, const float* res_f32, const bf16* aux, float* out_f32,
                bf16* out_bf16, cudaStream_t st, bool splitk = false, float alpha = 1.f);
  int linear_dwb(const bf16* dY, int M, const LinW& w, const bf16* X, cudaStream_t st, bool shared_weight = false);
  // drop = dropout site of the attention weights (0 = none); P2 receives dropout(P) and multiplies V
  int attn_fwd(const bf16* Q, long long ldq, const bf16* K, long long ldk, const bf16* V, long long ldv, const uint8_t* mask,
               bf16* P, bf16* O, int B, int Lq, int Lk, cudaStream_t st, bool v_on_aux = false, float* S = nullptr,
               uint32_t drop = 0, bf16* P2 = nullptr);
  int attn_bwd(const bf16* dO, const bf16* Q, long long ldq, const bf16* K, long long ldk, const bf16* V, long long ldv,
               const bf16* P, bf16* dQ, long long lddq, bf16* dK, long long lddk, bf16* dV, long long lddv, int B, int Lq,
               int Lk, cudaStream_t st, float* S = nullptr, bf16* dS = nullptr, bool use_aux = true, uint32_t drop = 0,
               const bf16* P2 = nullptr, const bf16* O = nullptr);   // O: the forward's attention output [B*Lq, 512] (optional)
  int enc_layer_fwd(const EncW& w, EncActs& a, const float* x, const float* pos, const float* next_pos, int L,
                    const uint8_t* mask, int B, bool emit_next, bf16* next_xb, bf16* next_xpb, cudaStream_t st, uint32_t site);
  int enc_layer_bwd(const EncW& w, EncActs& a, const float* x_in, const float* dx2, const float* dx2b, float* dx_out,
                    float* gxp_out, int L, int B, cudaStream_t st, uint32_t site);
  int backbone_fwd(int cam, int B, const float* image, cudaStream_t st);
  int backbone_bwd(int cam, int B, cudaStream_t st);
  int dec_self_fwd(int B, cudaStream_t st);     // query self-attention block: depends on weights only
  int dec_cross_fwd(int B, cudaStream_t st);    // cross-attention + FFN + norms on the encoder memory
  int transformer_bwd(float loss_grad, cudaStream_t st);
  int cvae_fwd(int B, const float* qpos, const float* actions, long long actions_bs, const uint8_t* is_pad,
               long long is_pad_bs, const float* eps, cudaStream_t st);
  int cvae_bwd(cudaStream_t st);

  template <typename T> T* alloc(long long n);
  TensorEntry* find(const std::string& name);
  float* P(const std::string& name) { return params_ + find(name)->offset; }
  float* G(const std::string& name) { return grads_ ? grads_ + find(name)->offset : nullptr; }
  const bf16* S(const std::string& name) { return shadow_ + find(name)->offset; }
  LinW lin(const std::string& wname, const std::string& bname, int K, int row0 = 0, int rows = -1);
  ConvW convw(const std::string& wname, const std::string& bnprefix, int Cin, int Cout, int KH, int stride, int pad);
  LnW lnw(const std::string& prefix);
  void build_tensor_table();
  void setup_views();
  int alloc_workspace();

  EngineConfig cfg_;
  std::vector<TensorEntry> tensors_;
  std::map<std::string, int> index_;
  long long total_floats_ = 0;
  long long region_off_[R_COUNT] = {0}, region_n_[R_COUNT] = {0};
  float *params_ = nullptr, *grads_ = nullptr, *m_ = nullptr, *v_ = nullptr;
  bf16* shadow_ = nullptr;
  DpPeers peers_;
  uint32_t *dp_signal_ = nullptr, *dp_epoch_ = nullptr, *dp_status_ = nullptr;
  cudaStream_t dp_stream_ = nullptr, dp_pack_stream_ = nullptr;
  cudaEvent_t dp_pack_ev_ = nullptr, dp_pack_done_ = nullptr;
  bool dp_pack_running_ = false;
  cudaEvent_t dp_ev_[DP_MAX_SITES] = {}, dp_done_ = nullptr;
  int dp_site_ = 0;
  bool dp_pack_pending_ = false;                // backbone updated by the p2p step: re-pack after the closing barrier
  int repack_conv_operands(cudaStream_t st, int cam = -1, bool stem = true);   // cam < 0: every camera
  int repack_pair_operands(int cam, cudaStream_t st);   // after the packed conv operands of that camera were rebuilt
  bool dp_fused_ = false;                       // true while backward_and_step(p2p = true) is being enqueued
  std::vector<void*> dp_opened_;
  std::vector<std::pair<long long, long long>> dp_ranges_;   // ranges exchanged by the last p2p step
  // peer-memory step: the ranges that become final last (backbone, CVAE head / tail) are not exchanged one by one with a
  // barrier each, but collected here and exchanged behind ONE barrier by kernels running side by side (dp_flush_deferred)
  struct DpSeg { AdamWArgs a; long long off; cudaStream_t producer; bool repack; };
  std::vector<DpSeg> dp_deferred_;
  bool dp_defer_ = false, dp_defer_now_ = false;   // policy of this step / active for the range being enqueued
  cudaStream_t dp_aux_[3] = {};
  cudaEvent_t dp_aux_ev_[8] = {};
  int dp_flush_deferred();
  int dp_sync_point(cudaStream_t producer);     // dp_stream_ waits for `producer`, then a cross-rank flag barrier on dp_stream_
  std::vector<void*> allocs_;
  long long ws_bytes_ = 0;
  int launches_ = 0;
  // Independent sub-graphs (CVAE encoder vs. camera backbones, forward and backward) run on a side stream forked from /
  // joined into the caller's stream with events; under stream capture this becomes two parallel branches of the graph.
  cudaStream_t side_ = nullptr;
  cudaEvent_t ev_fork_ = nullptr, ev_join_ = nullptr;
  bool concurrent_ = true;                      // ACT_CONCURRENT=0 serialises everything on the caller's stream
  int sm_reserve_ = 0;                          // ACT_SM_RESERVE: SMs the backbone launches leave to the concurrent CVAE branch
  int fork_side(cudaStream_t st);               // side_ waits for everything enqueued on st so far
  int join_side(cudaStream_t st);               // st waits for everything enqueued on side_ so far
  // Weight/bias gradients are leaves of the backward graph (nothing reads them before the optimiser), so they are
  // issued on a per-branch weight-gradient stream that trails the dX chain and fills the SMs its small grids leave idle.
  // Their dY operands come from a bump arena (gy_alloc) instead of reused scratch, so the dX chain never has to wait
  // for a weight-gradient read before overwriting a buffer.
  static constexpr int NWG = 3;                 // weight gradients of different layers are independent: round-robin
  cudaStream_t wg_[2][NWG + 1] = {};            // [0][*] trail the caller's stream, [1][*] trail side_; lane NWG = optimiser
  cudaEvent_t wg_ev_[2] = {nullptr, nullptr}, wg_done_[2][NWG + 1] = {}, ev_stage_ = nullptr;
  bool wg_used_[2][NWG + 1] = {};
  int wg_rr_[2] = {0, 0};
  int ensure_streams();
  cudaStream_t wg_fork(cudaStream_t st, int lane = -1);   // stream to launch a weight gradient on (st itself if serial)
  cudaStream_t wg_gather(cudaStream_t st, int lane = 0);  // that lane of st's branch, ordered after st and every other lane
  int wg_join(cudaStream_t st);                 // st waits for every weight-gradient launch issued so far
  // One auxiliary lane per branch for independent kernels inside a layer (V projection beside QK projection -> QK^T ->
  // softmax; dV / dK beside the dP -> softmax-backward -> dQ chain; dx(v) beside dx(qk)).
  // Lane 1 carries the decoder's query self-attention block (forward: weights only, so it runs beside the main encoder;
  // backward: a leaf after the cross-attention backward, so it runs beside the encoder backward) on private scratch.
  cudaStream_t aux_[2][2] = {};
  cudaEvent_t aux_ev_[2][2] = {}, aux_done_[2][2] = {};
  bool aux_used_[2][2] = {};
  cudaStream_t aux_fork(cudaStream_t st, int lane = 0);   // aux lane of st's branch, ordered after everything on st so far
  int aux_join(cudaStream_t st, int lane = 0);            // st waits for that aux lane
  float *Sbuf2_ = nullptr, *dgA_ = nullptr, *dgB_ = nullptr, *dgE_ = nullptr;
  bf16 *dSbuf2_ = nullptr, *dgb2_ = nullptr;
  // Long-K GEMMs with few output tiles (K = 3200: FFN down-projection forward, FFN up-projection data gradient) run
  // split-K with fp32 reductions into PRE-ZEROED outputs (bias / residual are added by the first split only).  The
  // zeroing happens on a weight-gradient lane at the start of forward / backward, long before the outputs are needed.
  cudaEvent_t ev_zero_f_ = nullptr, ev_zero_b_ = nullptr;
  bool zero_f_pending_ = false, zero_b_pending_ = false;
  float* gz_arena_ = nullptr;                   // zeroed fp32 outputs of the backward split-K GEMMs (one slice per layer)
  long long gz_cap_ = 0, gz_off_ = 0;
  float* gz_alloc(long long elems);
  long long gz_cvae_base_ = 0;
  static bool splitk_on() { static int v = -1; if (v < 0) { const char* e = getenv("ACT_SPLITK"); v = !(e && e[0] == '0'); } return v != 0; }
  // the forward stays bit-reproducible by default (split-K reduction order is not); ACT_SPLITK_FWD=1 opts in
  static bool splitk_fwd_on() { static int v = -1; if (v < 0) { const char* e = getenv("ACT_SPLITK_FWD"); v = (e && e[0] == '1'); } return v != 0 && splitk_on(); }
  int zero_backward_outputs(cudaStream_t st);
  int zero_forward_outputs(cudaStream_t st, bool with_cvae);
  bf16* gy_arena_ = nullptr;
  long long gy_cap_ = 0, gy_off_ = 0, gy_lim_ = 0, gy_cvae_base_ = 0;
  bf16* gy_alloc(long long elems);
  bf16* cgy_arena_ = nullptr;                   // same idea for the backbone's activation gradients (conv wgrad operands)
  long long cgy_cap_ = 0, cgy_off_ = 0;
  bf16* cgy_alloc(long long elems);

  // geometry
  int Hf_ = 0, Wf_ = 0, L_ = 0, Lp_ = 0, Lc_ = 0, Lcp_ = 0, nqp_ = 0;
  int fh_[6] = {0}, fw_[6] = {0};   // spatial dims: [0] image, [1] conv1 out, [2] pool out (layer1), [3] layer2, [4] layer3, [5] layer4
  int cur_B_ = 0;
  bool have_actions_ = false;
  float drop_p_ = 0.f;                          // configured dropout probability (0 = off)
  unsigned long long drop_seed_ = 0;
  uint32_t* drop_step_dev_ = nullptr;           // device step counter
  bool drop_active_ = false;                    // the last forward applied dropout (its backward regenerates the masks)

  // weights
  std::vector<EncW> enc_w_, cvae_w_;
  std::vector<DecW> dec_w_;
  LnW dec_norm_;
  std::vector<CamW> cam_w_;
  LinW input_proj_;

  // workspace
  std::vector<EncActs> enc_a_, cvae_a_;
  DecActs dec_a_;
  std::vector<CamActs> cam_a_;
  int Wp_ = 0;                  // padded row length (pixels) of the stem input
  float *tokC_ = nullptr, *src_ = nullptr, *pos_main_ = nullptr, *tgt0_ = nullptr;
  bf16 *mem_b_ = nullptr, *mem_pb_ = nullptr;
  uint8_t* maskC_ = nullptr;
  float *Sbuf_ = nullptr; bf16* dSbuf_ = nullptr;
  float *latent_info_ = nullptr, *mu_ = nullptr, *logvar_ = nullptr, *z_ = nullptr, *eps_ = nullptr;
  float *hs_ = nullptr, *hs_mean_ = nullptr, *hs_rstd_ = nullptr, *a_hat_ = nullptr, *is_pad_hat_ = nullptr, *losses_ = nullptr;
  float* one_ = nullptr;
  // saved input pointers of the last forward (for the backward)
  const float *in_qpos_ = nullptr, *in_image_ = nullptr, *in_actions_ = nullptr;
  long long in_actions_bs_ = 0;
  const uint8_t* in_is_pad_ = nullptr; long long in_is_pad_bs_ = 0;
  // backward scratch
  float *d_a_hat_ = nullptr, *d_mu_ = nullptr, *d_logvar_ = nullptr, *d_hs_ = nullptr, *dz_ = nullptr, *d_info_ = nullptr;
  float *gA_ = nullptr, *gB_ = nullptr, *gC_ = nullptr, *gD_ = nullptr, *gE_ = nullptr;   // fp32 [Mmax, 512] ping-pong gradient buffers
  float *gMem_ = nullptr, *gMemPos_ = nullptr;
  bf16 *gb1_ = nullptr, *gb2_ = nullptr, *gbqk_ = nullptr, *gbhh_ = nullptr, *gbv_ = nullptr, *gbk_ = nullptr;
  float* d_src_ = nullptr; float* d_tokC_ = nullptr;
  bf16 *cg_[4] = {nullptr, nullptr, nullptr, nullptr};   // backbone gradient ping-pong (NHWC bf16, max size)
  bf16* d_feat_ = nullptr;      // [K][B*Hf*Wf*512] gradient wrt each camera's layer4 output (ReLU-masked)
  float* wstage_ = nullptr; long long wstage_n_ = 0;   // staged conv weight gradients T[KK][Cout][Cin] of one backbone
  float* conv1_tmp_ = nullptr;  // [64][64][7] accumulator of the stem weight gradient (window-element layout)
  int stem_launch(int cam, int B, ConvLaunch* c, WgradLaunch* w);
  int input_grads(int B, cudaStream_t st);
};

}  // namespace act
