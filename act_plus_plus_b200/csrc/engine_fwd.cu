// Engine: workspace, forward pass (CVAE encoder, ResNet-18 backbones, transformer encoder/decoder, heads, loss).
#include "engine.h"
#include <cmath>
#include <cstring>

namespace act {

#define RUN(expr) do { ++launches_; if ((expr) != 0) return -1; } while (0)

static int roundup(int a, int b) { return (a + b - 1) / b * b; }

int Engine::init() {
  ACT_CHECK(cfg_.hidden_dim == 512 && cfg_.nheads == 8, "engine kernels are specialised for hidden_dim 512 / 8 heads (head_dim 64)");
  ACT_CHECK(cfg_.dim_ff % 64 == 0, "dim_feedforward must be a multiple of 64");
  ACT_CHECK(cfg_.num_cameras >= 1 && cfg_.num_cameras <= 8 && cfg_.batch_max >= 1, "bad camera / batch count");
  ACT_CHECK(cfg_.action_dim <= 128 && cfg_.state_dim <= 128 && cfg_.latent_dim <= 128, "small dims must be <= 128");
  fh_[0] = cfg_.img_h; fw_[0] = cfg_.img_w;
  fh_[1] = conv_out_dim(fh_[0], 7, 2, 3); fw_[1] = conv_out_dim(fw_[0], 7, 2, 3);
  fh_[2] = conv_out_dim(fh_[1], 3, 2, 1); fw_[2] = conv_out_dim(fw_[1], 3, 2, 1);
  for (int i = 3; i <= 5; ++i) {
    ACT_CHECK(fh_[i - 1] % 2 == 0 && fw_[i - 1] % 2 == 0, "image size must keep every stride-2 stage input even (e.g. multiples of 32)");
    fh_[i] = conv_out_dim(fh_[i - 1], 3, 2, 1); fw_[i] = conv_out_dim(fw_[i - 1], 3, 2, 1);
  }
  Hf_ = fh_[5]; Wf_ = fw_[5];
  L_ = 2 + Hf_ * Wf_ * cfg_.num_cameras; Lp_ = roundup(L_, 8);
  Lc_ = cfg_.num_queries + 2; Lcp_ = roundup(Lc_, 8);
  nqp_ = roundup(cfg_.num_queries, 8);
  ACT_CHECK(Lp_ <= 1024, "too many encoder tokens (max 1024)");
  build_tensor_table();
  cam_w_.resize(cfg_.num_cameras);
  return alloc_workspace();
}

int Engine::alloc_workspace() {
  const int Bm = cfg_.batch_max, d = cfg_.hidden_dim, ff = cfg_.dim_ff, h = cfg_.nheads, nq = cfg_.num_queries;
  const long long Me = 1LL * Bm * L_, Mc = 1LL * Bm * Lc_, Md = 1LL * Bm * nq;
  const long long Mmax = std::max(Me, std::max(Mc, Md));
#define A_(T, n) ({ T* _p = alloc<T>(n); if (_p == nullptr) return -1; _p; })
  shadow_ = A_(bf16, total_floats_);
  one_ = A_(float, 8);
  drop_step_dev_ = reinterpret_cast<uint32_t*>(A_(float, 8));
  opt_step_dev_ = reinterpret_cast<int*>(A_(float, 8));
  qnorm_mean_ = A_(float, 128); qnorm_std_ = A_(float, 128);
  { float one = 1.f; cudaMemcpy(one_, &one, sizeof(float), cudaMemcpyHostToDevice); }
  // packed conv operands + folded BN
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    CamW& c = cam_w_[k];
    c.stem.wpk = A_(bf16, 256 * 448); c.stem.scale = A_(float, 256); c.stem.shift = A_(float, 256);   // grouped layout, see pack_stem_kernel
    static const int cin_[4] = {64, 64, 128, 256}, cout_[4] = {64, 128, 256, 512};
    for (int s = 0; s < 4; ++s)
      for (int b = 0; b < 2; ++b) {
        BlockW& bw = c.blk[s * 2 + b];
        const int cin = b == 0 ? cin_[s] : cout_[s], cout = cout_[s];
        bw.c1.wpk = A_(bf16, 9LL * cin * cout); bw.c1.wpkt = A_(bf16, 9LL * cin * cout);
        bw.c1.scale = A_(float, cout); bw.c1.shift = A_(float, cout);
        bw.c2.wpk = A_(bf16, 9LL * cout * cout); bw.c2.wpkt = A_(bf16, 9LL * cout * cout);
        bw.c2.scale = A_(float, cout); bw.c2.shift = A_(float, cout);
        if (cin == 64 && cout == 64 && conv_pair_ok(ConvShape{1, 2, 2, 64, 64, 3, 1, 1})) {   // ACT_CONV_PAIR=0 disables
          for (ConvW* w : {&bw.c1, &bw.c2}) {
            w->wpk2 = A_(bf16, 128 * 1152); w->wpkt2 = A_(bf16, 128 * 1152);
            w->scale2 = A_(float, 128); w->shift2 = A_(float, 128);
          }
        }
        if (b == 0 && s > 0) {
          bw.ds.wpk = A_(bf16, 1LL * cin * cout); bw.ds.wpkt = A_(bf16, 1LL * cin * cout);
          bw.ds.scale = A_(float, cout); bw.ds.shift = A_(float, cout);
        }
      }
  }
  // backbone activations
  cam_a_.resize(cfg_.num_cameras);
  Wp_ = roundup(2 * (fw_[1] - 1) + 16, 2);                 // last output pixel reads padded pixels [2(Wo-1), 2(Wo-1)+16)
  static const int cout_[4] = {64, 128, 256, 512};
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    CamActs& a = cam_a_[k];
    a.xp = A_(bf16, 1LL * Bm * fh_[0] * Wp_ * 4 + 64);
    a.y1 = A_(bf16, 1LL * Bm * fh_[1] * fw_[1] * 64);
    a.p1 = A_(bf16, 1LL * Bm * fh_[2] * fw_[2] * 64);
    a.pidx = A_(uint8_t, 1LL * Bm * fh_[2] * fw_[2] * 64);
    for (int s = 0; s < 4; ++s)
      for (int b = 0; b < 2; ++b) {
        const long long n = 1LL * Bm * fh_[2 + s] * fw_[2 + s] * cout_[s];
        a.blk[s * 2 + b].a = A_(bf16, n);
        a.blk[s * 2 + b].y = A_(bf16, n);
        if (b == 0 && s > 0) { a.blk[s * 2 + b].ds = A_(bf16, n); }
      }
  }
  // transformer activations
  auto enc_alloc = [&](EncActs& a, long long M, int L, int Lp) -> int {
    a.xb = A_(bf16, M * d); a.xpb = A_(bf16, M * d); a.qk = A_(bf16, M * 2 * d); a.v = A_(bf16, M * d);
    a.P = A_(bf16, 1LL * Bm * h * L * Lp); a.o = A_(bf16, M * d); a.x1b = A_(bf16, M * d); a.hh = A_(bf16, M * ff);
    if (cfg_.train) a.P2 = A_(bf16, 1LL * Bm * h * L * Lp);
    a.t1 = A_(float, M * d); a.x1 = A_(float, M * d); a.t2 = A_(float, M * d); a.x2 = A_(float, M * d);
    a.mean1 = A_(float, M); a.rstd1 = A_(float, M); a.mean2 = A_(float, M); a.rstd2 = A_(float, M);
    return 0;
  };
  enc_a_.resize(cfg_.enc_layers); cvae_a_.resize(cfg_.enc_layers);
  for (auto& a : enc_a_) if (enc_alloc(a, Me, L_, Lp_)) return -1;
  if (cfg_.train && !cfg_.no_encoder) for (auto& a : cvae_a_) if (enc_alloc(a, Mc, Lc_, Lcp_)) return -1;
  {
    DecActs& a = dec_a_;
    a.xb = A_(bf16, Md * d); a.xpb = A_(bf16, Md * d); a.qk = A_(bf16, Md * 2 * d); a.v = A_(bf16, Md * d);
    a.P = A_(bf16, 1LL * Bm * h * nq * nqp_); a.o = A_(bf16, Md * d);
    if (cfg_.train) { a.P2 = A_(bf16, 1LL * Bm * h * nq * nqp_); a.cP2 = A_(bf16, 1LL * Bm * h * nq * Lp_); }
    a.t1 = A_(float, Md * d); a.y1 = A_(float, Md * d); a.y1pb = A_(bf16, Md * d);
    a.mean1 = A_(float, Md); a.rstd1 = A_(float, Md);
    a.q = A_(bf16, Md * d); a.k = A_(bf16, Me * d); a.cv = A_(bf16, Me * d);
    a.cP = A_(bf16, 1LL * Bm * h * nq * Lp_); a.co = A_(bf16, Md * d);
    a.t2 = A_(float, Md * d); a.y2 = A_(float, Md * d); a.y2b = A_(bf16, Md * d);
    a.mean2 = A_(float, Md); a.rstd2 = A_(float, Md);
    a.hh = A_(bf16, Md * ff); a.t3 = A_(float, Md * d); a.y3 = A_(float, Md * d);
    a.mean3 = A_(float, Md); a.rstd3 = A_(float, Md);
  }
  tokC_ = A_(float, Mc * d); src_ = A_(float, Me * d); tgt0_ = A_(float, Md * d);
  pos_main_ = A_(float, 1LL * L_ * d);
  mem_b_ = A_(bf16, Me * d); mem_pb_ = A_(bf16, Me * d);
  maskC_ = A_(uint8_t, Mc);
  const long long smax = 1LL * Bm * h * std::max(1LL * L_ * Lp_, 1LL * Lc_ * Lcp_);
  Sbuf_ = A_(float, smax);
  Sbuf2_ = A_(float, 1LL * Bm * h * nq * nqp_);       // decoder self-attention scores (runs beside the encoder)
  latent_info_ = A_(float, 1LL * Bm * 2 * cfg_.latent_dim); mu_ = A_(float, 1LL * Bm * cfg_.latent_dim);
  logvar_ = A_(float, 1LL * Bm * cfg_.latent_dim); z_ = A_(float, 1LL * Bm * cfg_.latent_dim); eps_ = A_(float, 1LL * Bm * cfg_.latent_dim);
  hs_ = A_(float, Md * d); hs_mean_ = A_(float, Md); hs_rstd_ = A_(float, Md);
  a_hat_ = A_(float, Md * cfg_.action_dim); is_pad_hat_ = A_(float, Md); losses_ = A_(float, 8);
  if (cfg_.train) {
    dSbuf_ = A_(bf16, smax);
    dSbuf2_ = A_(bf16, 1LL * Bm * h * nq * nqp_);
    dgA_ = A_(float, Md * d); dgB_ = A_(float, Md * d); dgE_ = A_(float, Md * d); dgb2_ = A_(bf16, Md * d);
    d_a_hat_ = A_(float, Md * cfg_.action_dim); d_mu_ = A_(float, 1LL * Bm * cfg_.latent_dim);
    d_logvar_ = A_(float, 1LL * Bm * cfg_.latent_dim); d_hs_ = A_(float, Md * d); dz_ = A_(float, 1LL * Bm * cfg_.latent_dim);
    d_info_ = A_(float, 1LL * Bm * 2 * cfg_.latent_dim);
    gA_ = A_(float, Mmax * d); gB_ = A_(float, Mmax * d); gC_ = A_(float, Mmax * d); gD_ = A_(float, Mmax * d); gE_ = A_(float, Mmax * d);
    gMem_ = A_(float, Me * d); gMemPos_ = A_(float, Me * d);
    gb1_ = A_(bf16, Mmax * d); gb2_ = A_(bf16, Mmax * d); gbqk_ = A_(bf16, Mmax * 2 * d); gbhh_ = A_(bf16, Mmax * ff);
    gbv_ = A_(bf16, Mmax * d); gbk_ = A_(bf16, Mmax * d);
    {   // dY arena: every dY that feeds a deferred weight gradient (5 per encoder layer, 10 in the decoder layer, 1 input_proj)
      const long long per_row = 3LL * d + 2 * d + ff;                 // gb1 x2, gbv, gbqk, gbhh
      gy_cap_ = cfg_.enc_layers * (Me + Mc) * (per_row + 512) + (Md * (per_row + 4 * d) + 2 * Me * d) + 1LL * Bm * Hf_ * Wf_ * d * cfg_.num_cameras + 64 * 256;
      gy_arena_ = A_(bf16, gy_cap_);
      gz_cap_ = cfg_.enc_layers * (Me + Mc) * d + Md * d + 64 * 32;
      gz_arena_ = A_(float, gz_cap_);
      gz_cvae_base_ = gz_cap_ - (cfg_.enc_layers * Mc * d + 32 * 32);
      gy_cvae_base_ = gy_cap_ - (cfg_.enc_layers * Mc * (per_row + 512) + 16 * 256);
    }
    d_src_ = A_(float, Me * d); d_tokC_ = A_(float, Mc * d);
    const long long cmax = 1LL * Bm * fh_[1] * fw_[1] * 64;        // largest NHWC activation (conv1 output)
    for (int i = 0; i < 4; ++i) { cg_[i] = A_(bf16, i == 0 ? cmax : cmax / 4); }
    d_feat_ = A_(bf16, 1LL * cfg_.num_cameras * Bm * Hf_ * Wf_ * 512);
    {   // conv-gradient arena: per residual block the c2 input gradient and the block input gradient, + the pool gradient
      long long n = 1LL * Bm * fh_[1] * fw_[1] * 64 + 128;
      for (int i = 0; i < 8; ++i) {
        const int s4 = i / 2, Cout = 64 << s4, strided = (i % 2 == 0 && s4 > 0);
        const int Cin = strided ? Cout / 2 : Cout;
        n += 1LL * Bm * fh_[2 + s4] * fw_[2 + s4] * Cout + 1LL * Bm * fh_[strided ? 1 + s4 : 2 + s4] * fw_[strided ? 1 + s4 : 2 + s4] * Cin + 256;
      }
      cgy_cap_ = n * cfg_.num_cameras;
      cgy_arena_ = A_(bf16, cgy_cap_);
    }
    conv1_tmp_ = A_(float, 64 * 64 * 7);
    {
      static const int ci_[4] = {64, 64, 128, 256}, co_[4] = {64, 128, 256, 512};
      wstage_n_ = 0;
      for (int s4 = 0; s4 < 4; ++s4) {
        wstage_n_ += 9LL * ci_[s4] * co_[s4] + 3 * 9LL * co_[s4] * co_[s4];     // b0.c1, b0.c2, b1.c1, b1.c2
        if (s4 > 0) wstage_n_ += 1LL * ci_[s4] * co_[s4];                       // downsample 1x1
      }
      wstage_ = A_(float, wstage_n_);
    }
  }
#undef A_
  // static part of the encoder position table: rows 2.. = PositionEmbeddingSine (position_encoding.py:30-52,
  // normalize=True, scale 2*pi, temperature 1e4), cameras concatenated along W (detr_vae.py:217)
  {
    const int npf = d / 2;
    std::vector<float> pos(1LL * L_ * d, 0.f);
    const float two_pi = 2.0f * 3.14159265358979323846f;
    for (int y = 0; y < Hf_; ++y)
      for (int k = 0; k < cfg_.num_cameras; ++k)
        for (int x = 0; x < Wf_; ++x) {
          const float ye = static_cast<float>(y + 1) / (static_cast<float>(Hf_) + 1e-6f) * two_pi;
          const float xe = static_cast<float>(x + 1) / (static_cast<float>(Wf_) + 1e-6f) * two_pi;
          float* row = pos.data() + (2LL + (1LL * y * cfg_.num_cameras + k) * Wf_ + x) * d;
          for (int i = 0; i < npf; ++i) {
            const float dim_t = std::pow(10000.0f, 2.0f * static_cast<float>(i / 2) / static_cast<float>(npf));
            const float py = ye / dim_t, px = xe / dim_t;
            row[i] = (i % 2 == 0) ? std::sin(py) : std::cos(py);
            row[npf + i] = (i % 2 == 0) ? std::sin(px) : std::cos(px);
          }
        }
    ACT_CUDA(cudaMemcpy(pos_main_, pos.data(), pos.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  return 0;
}

// --------------------------------------------------------------------------------------------- GEMM helpers
int Engine::linear(const bf16* X, int M, const LinW& w, uint32_t flags, const float* res_f32, float* out_f32,
                   bf16* out_bf16, const bf16* aux, cudaStream_t st, bool splitk, uint32_t drop) {
  MatDesc A, B;
  A.ptr = X; A.ld = w.K; B.ptr = w.w; B.ld = w.K;
  EpiDesc e;
  e.flags = flags | (w.b ? EPI_BIAS : 0) | (out_f32 ? EPI_OUT_F32 : 0) | (out_bf16 ? EPI_OUT_BF16 : 0) |
            (res_f32 ? EPI_RES_F32 : 0) | (aux ? EPI_MASK_POS : 0);
  if (drop && drop_active_) { e.flags |= EPI_DROPOUT; e.drop = drop_site(drop); splitk = false; }
  e.bias = w.b; e.res_f32 = res_f32; e.out_f32 = out_f32; e.out_bf16 = out_bf16; e.aux_bf16 = aux;
  e.ldc = w.N; e.ldr = w.N;
  int bn = 0, ks = 1;
  // opt-in, training calls only: the fp32 reduction order of a split-K launch is not reproducible
  if (splitk && splitk_fwd_on() && have_actions_ && out_f32 && !out_bf16 && !aux && !(flags & EPI_RELU)) {   // caller pre-zeroed out_f32
    plan_splitk(M, w.N, w.K, &bn, &ks);
    if (ks > 1) e.flags |= EPI_ATOMIC; else bn = 0;
  }
  ++launches_;
  return tc_gemm(M, w.N, w.K, 1, 1, A, B, e, ks, st, bn);
}

// Zero the fp32 outputs of the forward split-K GEMMs (the FFN down-projections) on a weight-gradient lane.
int Engine::zero_forward_outputs(cudaStream_t st, bool with_cvae) {
  zero_f_pending_ = false;
  if (!splitk_fwd_on() || !have_actions_) return 0;
  cudaStream_t z = wg_fork(st, 1);
  const int d = cfg_.hidden_dim;
  const long long Me = 1LL * cur_B_ * L_, Mc = 1LL * cur_B_ * Lc_, Md = 1LL * cur_B_ * cfg_.num_queries;
  for (auto& a : enc_a_) ACT_CUDA(cudaMemsetAsync(a.t2, 0, sizeof(float) * Me * d, z));
  if (with_cvae) for (auto& a : cvae_a_) ACT_CUDA(cudaMemsetAsync(a.t2, 0, sizeof(float) * Mc * d, z));
  ACT_CUDA(cudaMemsetAsync(dec_a_.t3, 0, sizeof(float) * Md * d, z));
  zero_f_pending_ = z != st;
  if (zero_f_pending_) ACT_CUDA(cudaEventRecord(ev_zero_f_, z));
  return 0;
}

// softmax(Q K^T / sqrt(64) + mask) V per (batch, head); Q/K/V are strided head views of projection buffers.
int Engine::attn_fwd(const bf16* Q, long long ldq, const bf16* K, long long ldk, const bf16* V, long long ldv,
                     const uint8_t* mask, bf16* P, bf16* O, int B, int Lq, int Lk, cudaStream_t st, bool v_on_aux, float* S,
                     uint32_t drop, bf16* P2) {
  if (!S) S = Sbuf_;
  const bool dropping = drop && drop_active_;
  ACT_CHECK(!dropping || P2, "attention dropout needs the second probability buffer (train workspace)");
  const int h = cfg_.nheads, hd = 64, Lkp = roundup(Lk, 8);
  MatDesc a, b;
  EpiDesc e;
  a.ptr = Q; a.ld = ldq; a.sz0 = hd; a.sz1 = Lq * ldq;
  b.ptr = K; b.ld = ldk; b.sz0 = hd; b.sz1 = Lk * ldk;
  if (tc_attn_scores_applicable(Lk, Lkp, dropping)) {
    // scores + mask + softmax in one kernel: the fp32 score tile stays in TMEM (csrc/attn_kernel.cuh)
    ++launches_;
    const DropDesc dd = dropping ? drop_site(drop) : DropDesc();
    if (tc_attn_scores(0, a, b, Lq, Lk, Lkp, h, B, 0.125f, mask, nullptr, P, st, dropping ? &dd : nullptr, dropping ? P2 : nullptr)) return -1;
  } else {
    ACT_CHECK(!dropping, "attention dropout is implemented in the fused scores kernels only (ACT_FUSED_ATTN=0 is set)");
    e.flags = EPI_OUT_F32; e.alpha = 0.125f; e.out_f32 = S;
    e.ldc = Lkp; e.osz0 = 1LL * Lq * Lkp; e.osz1 = 1LL * h * Lq * Lkp;
    ++launches_;
    if (tc_gemm(Lq, Lk, hd, h, B, a, b, e, 1, st)) return -1;
    RUN(softmax_fwd(S, mask, P, B * h * Lq, Lk, Lkp, h * Lq, st));
  }
  if (v_on_aux && aux_join(st)) return -1;           // V was projected on the aux lane
  MatDesc pa, vb;
  EpiDesc eo;
  pa.ptr = dropping ? P2 : P; pa.ld = Lkp; pa.sz0 = 1LL * Lq * Lkp; pa.sz1 = 1LL * h * Lq * Lkp;
  vb.ptr = V; vb.mn_major = 1; vb.ld = ldv; vb.sz0 = hd; vb.sz1 = Lk * ldv;
  eo.flags = EPI_OUT_BF16; eo.out_bf16 = O; eo.ldc = 512; eo.osz0 = hd; eo.osz1 = 1LL * Lq * 512;
  ++launches_;
  return tc_gemm(Lq, hd, Lk, h, B, pa, vb, eo, 1, st);
}

// One post-norm encoder layer (transformer.py:211-224).  a.xb / a.xpb (bf16 x and x+pos) are filled by the caller;
// the second LayerNorm also emits the next consumer's bf16(x) / bf16(x + next_pos).
int Engine::enc_layer_fwd(const EncW& w, EncActs& a, const float* x, const float* pos, const float* next_pos, int L,
                          const uint8_t* mask, int B, bool emit_next, bf16* next_xb, bf16* next_xpb, cudaStream_t st, uint32_t site) {
  (void)pos;
  const int M = B * L, d = cfg_.hidden_dim;
  cudaStream_t ax = aux_fork(st);                    // V projection beside QK projection -> QK^T -> softmax
  if (linear(a.xb, M, w.sa.v, 0, nullptr, nullptr, a.v, nullptr, ax)) return -1;
  if (linear(a.xpb, M, w.sa.qk, 0, nullptr, nullptr, a.qk, nullptr, st)) return -1;
  if (attn_fwd(a.qk, 2 * d, a.qk + d, 2 * d, a.v, d, mask, a.P, a.o, B, L, L, st, ax != st, nullptr, site + 0, a.P2)) return -1;
  if (linear(a.o, M, w.sa.o, 0, x, a.t1, nullptr, nullptr, st, false, site + 1)) return -1;      // x + dropout1(attn)
  RUN(ln_fwd(a.t1, w.n1.g, w.n1.b, nullptr, 1, a.x1, a.x1b, nullptr, a.mean1, a.rstd1, M, d, st));
  if (linear(a.x1b, M, w.l1, EPI_RELU, nullptr, nullptr, a.hh, nullptr, st, false, site + 2)) return -1;   // dropout(relu(.))
  if (zero_f_pending_) ACT_CUDA(cudaStreamWaitEvent(st, ev_zero_f_, 0));      // a.t2 was zeroed for the split-K reduction
  if (linear(a.hh, M, w.l2, 0, a.x1, a.t2, nullptr, nullptr, st, true, site + 3)) return -1;     // x + dropout2(ffn)
  RUN(ln_fwd(a.t2, w.n2.g, w.n2.b, next_pos, L, a.x2, emit_next ? next_xb : nullptr, emit_next ? next_xpb : nullptr,
             a.mean2, a.rstd2, M, d, st));
  return 0;
}


// small kernels local to the forward -----------------------------------------------------------------
__global__ void build_cvae_mask_kernel(const uint8_t* __restrict__ is_pad, long long bs, uint8_t* __restrict__ mask, int B, int nq) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * (nq + 2)) return;
  const int b = i / (nq + 2), t = i % (nq + 2);
  mask[i] = t < 2 ? 0 : (is_pad[b * bs + (t - 2)] != 0);     // cls / qpos tokens are never padding (detr_vae.py:127-128)
}

// DETRVAE.encode training branch (detr_vae.py:115-151)
int Engine::cvae_fwd(int B, const float* qpos, const float* actions, long long actions_bs, const uint8_t* is_pad,
                     long long is_pad_bs, const float* eps, cudaStream_t st) {
  ACT_CHECK(cfg_.train && !cfg_.no_encoder, "this engine was created without the CVAE-encoder workspace (train = 0 or no_encoder)");
  const int d = cfg_.hidden_dim, nq = cfg_.num_queries, S = cfg_.state_dim, A = cfg_.action_dim, Z = cfg_.latent_dim;
  SgemmArgs g;
  // token 0: cls_embed broadcast over the batch
  g = SgemmArgs(); g.I = B; g.J = d; g.R = 1; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = P("cls_embed.weight"); g.br = act::lin(0); g.bj = act::lin(1); g.C = tokC_; g.ci = act::lin(1LL * Lc_ * d); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  // token 1: encoder_joint_proj(qpos)
  g = SgemmArgs(); g.I = B; g.J = d; g.R = S; g.A = qpos; g.ai = act::lin(S); g.ar = act::lin(1);
  g.B = P("encoder_joint_proj.weight"); g.br = act::lin(1); g.bj = act::lin(S); g.bias = P("encoder_joint_proj.bias");
  g.C = tokC_ + d; g.ci = act::lin(1LL * Lc_ * d); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  // tokens 2..: encoder_action_proj(actions[:, :nq])
  g = SgemmArgs(); g.I = B * nq; g.J = d; g.R = A; g.A = actions; g.ai = batched(nq, actions_bs, A); g.ar = act::lin(1);
  g.B = P("encoder_action_proj.weight"); g.br = act::lin(1); g.bj = act::lin(A); g.bias = P("encoder_action_proj.bias");
  g.C = tokC_ + 2 * d; g.ci = batched(nq, 1LL * Lc_ * d, d); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  build_cvae_mask_kernel<<<(B * Lc_ + 127) / 128, 128, 0, st>>>(is_pad, is_pad_bs, maskC_, B, nq);
  ++launches_;
  ACT_CUDA(cudaGetLastError());
  const float* pos = P("pos_table");
  RUN(add_pos_cast(tokC_, pos, Lc_, cvae_a_[0].xb, cvae_a_[0].xpb, B * Lc_, d, st));
  const int nl = cfg_.enc_layers;
  for (int l = 0; l < nl; ++l) {
    const float* x = l == 0 ? tokC_ : cvae_a_[l - 1].x2;
    const bool more = l + 1 < nl;
    if (enc_layer_fwd(cvae_w_[l], cvae_a_[l], x, pos, pos, Lc_, maskC_, B, more, more ? cvae_a_[l + 1].xb : nullptr,
                      more ? cvae_a_[l + 1].xpb : nullptr, st, SITE_CVAE + 8 * l)) return -1;
  }
  // latent_proj on the cls output rows
  g = SgemmArgs(); g.I = B; g.J = 2 * Z; g.R = d; g.A = cvae_a_[nl - 1].x2; g.ai = act::lin(1LL * Lc_ * d); g.ar = act::lin(1);
  g.B = P("latent_proj.weight"); g.br = act::lin(1); g.bj = act::lin(d); g.bias = P("latent_proj.bias");
  g.C = latent_info_; g.ci = act::lin(2 * Z); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  ACT_CUDA(cudaMemcpyAsync(eps_, eps, sizeof(float) * B * Z, cudaMemcpyDeviceToDevice, st));
  RUN(reparam_fwd(latent_info_, eps_, mu_, logvar_, z_, B, Z, st));
  return 0;
}

// Stem conv 7x7/s2/p3 (3 -> 64) as an implicit GEMM over the padded NHWC4 image: one tap per kernel row kh, whose
// A row for output pixel (h, w) is the 16-pixel x 4-channel window starting at padded pixel 2w of input row 2h+kh-3
// (a TMA dimension with a 16-byte stride; rows outside the image are TMA zero fill).  K = 7 x 64, 147 of them non-zero.
int Engine::stem_launch(int cam, int B, ConvLaunch* c, WgradLaunch* w) {
  ACT_CHECK(fh_[0] % 2 == 0, "stem needs an even image height");
  const CamW& cw = cam_w_[cam];
  const CamActs& ca = cam_a_[cam];
  const int H = fh_[0], Ho = fh_[1], Wo = fw_[1];
  PlaneDesc pl[2];
  for (int ph = 0; ph < 2; ++ph) {
    pl[ph].ptr = ca.xp + 1LL * ph * Wp_ * 4; pl[ph].C = 64; pl[ph].W = Wo; pl[ph].H = H / 2; pl[ph].N = B;
    pl[ph].sW = 8; pl[ph].sH = 2LL * Wp_ * 4; pl[ph].sN = 1LL * H * Wp_ * 4;
  }
  TapDesc taps[7];
  for (int kh = 0; kh < 7; ++kh) {
    const int oy = kh - 3;
    taps[kh].map = static_cast<int16_t>(((oy % 2) + 2) % 2);
    taps[kh].dh = static_cast<int16_t>(oy >= 0 ? oy / 2 : -((-oy + 1) / 2));
    taps[kh].dw = 0; taps[kh].wslot = static_cast<int16_t>(kh);
  }
  if (c) {
    // Forward: one GEMM row = the 16-pixel window starting at output pixel g*r and produces the g = 4 (or 2) adjacent
    // output pixels whose taps all lie inside it; N = 64*g = their channels one after the other (weights rows 64*j.. are
    // the filters shifted by 8*j columns).  The output [B, Ho, Wo, 64] is the same memory as [B, Ho, Wo/g, 64*g].
    // The issuer sustains one M=128 MMA per ~55 ns whatever N <= 128 is and ~75 ns at N = 256 (section 4.1 of DESIGN.md),
    // so wider N is the only way to fill the tensor pipe with a 64-channel output.  ACT_STEM_GROUP=1|2|4 overrides.
    const char* ge = getenv("ACT_STEM_GROUP");             // read per call: the parity tests switch it between engines
    const int grp_env = ge ? atoi(ge) : 4;
    int g = grp_env >= 4 ? 4 : grp_env >= 2 ? 2 : 1;
    while (g > 1 && Wo % g) g /= 2;
    c->planes[0] = pl[0]; c->planes[1] = pl[1]; c->nplanes = 2; c->ntaps = 7;
    for (int i = 0; i < 7; ++i) c->taps[i] = taps[i];
    c->Cin = 64; c->wgt = cw.stem.wpk; c->wgt_ld = 448;
    for (int ph = 0; ph < 2; ++ph) { c->planes[ph].W = Wo / g; c->planes[ph].sW = 8 * g; }
    c->Cout = 64 * g; c->vW = Wo / g; c->vH = Ho; c->NIMG = B;
    c->og = PixGeom{Ho, Wo / g, 1, 1, 0, 0, 64 * g, 0};
    c->rg = c->og;
    c->algo_flops = 2.0 * B * Ho * Wo * 64 * 147;           // 7x7x3 taps; the K-slabs are zero-padded to 7x64
  }
  if (w) {
    w->algo_flops = 2.0 * B * Ho * Wo * 64 * 147;
    w->planes[0] = pl[0]; w->planes[1] = pl[1]; w->nplanes = 2; w->ntaps = 7;
    for (int i = 0; i < 7; ++i) w->taps[i] = taps[i];
    w->Cin = 64; w->Cout = 64; w->KK = 7;
  }
  return 0;
}

// ResNet-18 trunk with FrozenBatchNorm folded into the conv epilogues (backbone.py:47-57,121-124), then the
// 1x1 input_proj (detr_vae.py:184) written straight into the encoder token buffer [B, L, d] at this camera's columns.
int Engine::backbone_fwd(int cam, int B, const float* image, cudaStream_t st) {
  const int K = cfg_.num_cameras, H = fh_[0], W = fw_[0], d = cfg_.hidden_dim;
  CamW& cw = cam_w_[cam];
  CamActs& ca = cam_a_[cam];
  if (image_fmt_ == 1) {
    const uint8_t* img8 = reinterpret_cast<const uint8_t*>(image) + 1LL * cam * H * W * 3;
    RUN(stem_pack_image_u8(img8, 1LL * K * H * W * 3, ca.xp, B, H, W, Wp_, st));
  } else {
    const float* img = image + 1LL * cam * 3 * H * W;
    RUN(stem_pack_image(img, 1LL * K * 3 * H * W, ca.xp, B, H, W, Wp_, st));
  }
  {
    ConvLaunch c;
    if (stem_launch(cam, B, &c, nullptr)) return -1;
    c.epi.flags = EPI_SCALE | EPI_BIAS | EPI_RELU | EPI_OUT_BF16; c.epi.scale = cw.stem.scale; c.epi.bias = cw.stem.shift;
    c.epi.out_bf16 = ca.y1;
    RUN(tc_conv(c, st));
  }
  RUN(maxpool_fwd(ca.y1, ca.p1, cfg_.train ? ca.pidx : nullptr, B, fh_[1], fw_[1], 64, st));
  const bf16* x = ca.p1;
  int Hc = fh_[2], Wc = fw_[2];
  for (int i = 0; i < 8; ++i) {
    BlockW& bw = cw.blk[i];
    BlockActs& ba = ca.blk[i];
    ConvShape s1{B, Hc, Wc, bw.c1.Cin, bw.c1.Cout, 3, bw.c1.stride, 1};
    EpiDesc e1;
    e1.flags = EPI_SCALE | EPI_BIAS | EPI_RELU | EPI_OUT_BF16; e1.scale = bw.c1.scale; e1.bias = bw.c1.shift; e1.out_bf16 = ba.a;
    const bf16* wp1 = (bw.c1.wpk2 && conv_pair_ok(s1)) ? bw.c1.wpk2 : nullptr;
    if (wp1) { e1.scale = bw.c1.scale2; e1.bias = bw.c1.shift2; }
    RUN(conv_fwd(s1, x, bw.c1.wpk, e1, st, wp1));
    const int Ho = conv_out_dim(Hc, 3, bw.c1.stride, 1), Wo = conv_out_dim(Wc, 3, bw.c1.stride, 1);
    const bf16* idt = x;
    if (bw.has_ds) {
      ConvShape sd{B, Hc, Wc, bw.ds.Cin, bw.ds.Cout, 1, bw.ds.stride, 0};
      EpiDesc ed;
      ed.flags = EPI_SCALE | EPI_BIAS | EPI_OUT_BF16; ed.scale = bw.ds.scale; ed.bias = bw.ds.shift; ed.out_bf16 = ba.ds;
      RUN(conv_fwd(sd, x, bw.ds.wpk, ed, st));
      idt = ba.ds;
    }
    ConvShape s2{B, Ho, Wo, bw.c2.Cin, bw.c2.Cout, 3, 1, 1};
    EpiDesc e2;
    e2.flags = EPI_SCALE | EPI_BIAS | EPI_RES_BF16 | EPI_RELU | EPI_OUT_BF16;
    e2.scale = bw.c2.scale; e2.bias = bw.c2.shift; e2.res_bf16 = idt; e2.out_bf16 = ba.y;
    const bf16* wp2 = (bw.c2.wpk2 && conv_pair_ok(s2)) ? bw.c2.wpk2 : nullptr;
    if (wp2) { e2.scale = bw.c2.scale2; e2.bias = bw.c2.shift2; }
    RUN(conv_fwd(s2, ba.a, bw.c2.wpk, e2, st, wp2));
    x = ba.y; Hc = Ho; Wc = Wo;
  }
  // input_proj: 1x1 conv 512 -> d with bias, fp32 rows (b, 2 + y*(Wf*K) + cam*Wf + x) of src_
  ConvLaunch c;
  c.planes[0].ptr = x; c.planes[0].C = 512; c.planes[0].W = Wf_; c.planes[0].H = Hf_; c.planes[0].N = B;
  c.planes[0].sW = 512; c.planes[0].sH = 512LL * Wf_; c.planes[0].sN = 512LL * Wf_ * Hf_;
  c.nplanes = 1; c.ntaps = 1; c.taps[0] = TapDesc{0, 0, 0, 0};
  c.Cin = 512; c.Cout = d; c.wgt = input_proj_.w; c.wgt_ld = 512;
  c.vW = Wf_; c.vH = Hf_; c.NIMG = B;
  c.og = PixGeom{Hf_, Wf_ * K, 1, 1, 0, cam * Wf_, d, 1LL * L_ * d};
  c.rg = c.og;
  c.epi.flags = EPI_BIAS | EPI_OUT_F32; c.epi.bias = input_proj_.b; c.epi.out_f32 = src_ + 2 * d;
  RUN(tc_conv(c, st));
  return 0;
}

// Decoder layer 0 (transformer.py:274-295) + the shared final norm (transformer.py:171-181); only index [0] of the
// stacked intermediates is consumed downstream (detr_vae.py:245).
int Engine::dec_self_fwd(int B, cudaStream_t st) {
  const int d = cfg_.hidden_dim, nq = cfg_.num_queries, Md = B * nq;
  const DecW& w = dec_w_[0];
  DecActs& a = dec_a_;
  const float* qpos_embed = P("query_embed.weight");
  RUN(add_pos_cast(tgt0_, qpos_embed, nq, a.xb, a.xpb, Md, d, st));            // tgt = 0 (transformer.py:109)
  if (linear(a.xpb, Md, w.sa.qk, 0, nullptr, nullptr, a.qk, nullptr, st)) return -1;
  if (linear(a.xb, Md, w.sa.v, 0, nullptr, nullptr, a.v, nullptr, st)) return -1;
  if (attn_fwd(a.qk, 2 * d, a.qk + d, 2 * d, a.v, d, nullptr, a.P, a.o, B, nq, nq, st, false, Sbuf2_, SITE_DEC + 0, a.P2)) return -1;
  if (linear(a.o, Md, w.sa.o, 0, tgt0_, a.t1, nullptr, nullptr, st, false, SITE_DEC + 1)) return -1;
  RUN(ln_fwd(a.t1, w.n1.g, w.n1.b, qpos_embed, nq, a.y1, nullptr, a.y1pb, a.mean1, a.rstd1, Md, d, st));
  if (linear(a.y1pb, Md, w.ca.q, 0, nullptr, nullptr, a.q, nullptr, st)) return -1;
  return 0;
}
int Engine::dec_cross_fwd(int B, cudaStream_t st) {
  const int d = cfg_.hidden_dim, nq = cfg_.num_queries, Md = B * nq, Me = B * L_;
  const DecW& w = dec_w_[0];
  DecActs& a = dec_a_;
  cudaStream_t ax = aux_fork(st);
  if (linear(mem_b_, Me, w.ca.v, 0, nullptr, nullptr, a.cv, nullptr, ax)) return -1;
  if (linear(mem_pb_, Me, w.ca.k, 0, nullptr, nullptr, a.k, nullptr, st)) return -1;
  if (attn_fwd(a.q, d, a.k, d, a.cv, d, nullptr, a.cP, a.co, B, nq, L_, st, ax != st, nullptr, SITE_DEC + 4, a.cP2)) return -1;
  if (linear(a.co, Md, w.ca.o, 0, a.y1, a.t2, nullptr, nullptr, st, false, SITE_DEC + 5)) return -1;
  RUN(ln_fwd(a.t2, w.n2.g, w.n2.b, nullptr, 1, a.y2, a.y2b, nullptr, a.mean2, a.rstd2, Md, d, st));
  if (linear(a.y2b, Md, w.l1, EPI_RELU, nullptr, nullptr, a.hh, nullptr, st, false, SITE_DEC + 2)) return -1;
  if (zero_f_pending_) ACT_CUDA(cudaStreamWaitEvent(st, ev_zero_f_, 0));
  if (linear(a.hh, Md, w.l2, 0, a.y2, a.t3, nullptr, nullptr, st, true, SITE_DEC + 3)) return -1;
  RUN(ln_fwd(a.t3, w.n3.g, w.n3.b, nullptr, 1, a.y3, nullptr, nullptr, a.mean3, a.rstd3, Md, d, st));
  RUN(ln_fwd(a.y3, dec_norm_.g, dec_norm_.b, nullptr, 1, hs_, nullptr, nullptr, hs_mean_, hs_rstd_, Md, d, st));
  return 0;
}

int Engine::set_qpos_norm(const float* mean_dev, const float* std_dev, cudaStream_t st) {
  ACT_CHECK(!mean_dev == !std_dev, "set_qpos_norm: mean and std must be given together");
  qnorm_on_ = mean_dev != nullptr;
  if (!qnorm_on_) return 0;
  qnorm_set_ = true;
  ACT_CUDA(cudaMemcpyAsync(qnorm_mean_, mean_dev, sizeof(float) * cfg_.state_dim, cudaMemcpyDeviceToDevice, st));
  ACT_CUDA(cudaMemcpyAsync(qnorm_std_, std_dev, sizeof(float) * cfg_.state_dim, cudaMemcpyDeviceToDevice, st));
  return 0;
}
int Engine::set_dropout(float p, unsigned long long seed) {
  ACT_CHECK(p >= 0.f && p < 1.f, "dropout probability must be in [0, 1)");
  drop_p_ = p; drop_seed_ = seed;
  return 0;
}
int Engine::dropout_state(float* p, unsigned long long* seed, uint32_t* step) {
  if (p) *p = drop_p_;
  if (seed) *seed = drop_seed_;
  if (step) ACT_CUDA(cudaMemcpy(step, drop_step_dev_, sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return 0;
}
DropDesc Engine::drop_site(uint32_t site) const {
  DropDesc d;
  if (!drop_active_) return d;
  d.seed_lo = static_cast<uint32_t>(drop_seed_); d.seed_hi = static_cast<uint32_t>(drop_seed_ >> 32);
  d.site = site; d.thresh = dropout_threshold(drop_p_); d.scale = 1.f / (1.f - drop_p_); d.step = drop_step_dev_;
  return d;
}

// Parity diagnostics (act_engine_debug_buffer): internal activations of the last forward.
int Engine::debug_buffer(const char* name, void** ptr, long long* numel, int* dtype) {
  ACT_CHECK(name && ptr && numel && dtype, "debug_buffer: null argument");
  ACT_CHECK(cur_B_ > 0, "debug_buffer: no forward has run");
  const std::string n(name);
  const long long d = cfg_.hidden_dim;
  if (n == "src") { *ptr = src_; *numel = 1LL * cur_B_ * L_ * d; *dtype = 0; return 0; }
  if (n == "memory") { *ptr = enc_a_[cfg_.enc_layers - 1].x2; *numel = 1LL * cur_B_ * L_ * d; *dtype = 0; return 0; }
  if (n == "hs") { *ptr = hs_; *numel = 1LL * cur_B_ * cfg_.num_queries * d; *dtype = 0; return 0; }
  if (n.compare(0, 4, "feat") == 0) {
    const int k = atoi(n.c_str() + 4);
    ACT_CHECK(k >= 0 && k < cfg_.num_cameras, "debug_buffer: camera index out of range");
    *ptr = cam_a_[k].blk[7].y; *numel = 1LL * cur_B_ * Hf_ * Wf_ * 512; *dtype = 1;
    return 0;
  }
  set_error("debug_buffer: unknown buffer " + n);
  return -1;
}

// ACTPolicy.__call__ (policy.py:264-332) minus the Python glue: normalise (fused into the stem's image pack), DETRVAE.forward
// (detr_vae.py:163-254), and for the training/validation call the L1 + KL loss (policy.py:309-318).
int Engine::forward(int B, const float* qpos, const float* image, const float* actions, long long actions_bs,
                    const uint8_t* is_pad, long long is_pad_bs, const float* eps, float* a_hat, float* is_pad_hat,
                    float* mu, float* logvar, float* losses, cudaStream_t st) {
  ACT_CHECK(params_ != nullptr, "engine not bound to parameters");
  ACT_CHECK(B >= 1 && B <= cfg_.batch_max, "batch exceeds the engine's batch_max");
  ACT_CHECK(qpos && image, "qpos / image must not be null");
  const int d = cfg_.hidden_dim, nq = cfg_.num_queries, S = cfg_.state_dim, A = cfg_.action_dim, Z = cfg_.latent_dim;
  cur_B_ = B; have_actions_ = actions != nullptr;
  // dropout applies to training calls of a train-mode policy (the reference's validation runs under .eval())
  drop_active_ = have_actions_ && cfg_.train && drop_p_ > 0.f;
  if (drop_active_) RUN(bump_counter(drop_step_dev_, st));
  in_qpos_ = qpos; in_image_ = image; in_actions_ = actions; in_actions_bs_ = actions_bs; in_is_pad_ = is_pad; in_is_pad_bs_ = is_pad_bs;
  ACT_CUDA(cudaMemcpyAsync(pos_main_, P("additional_pos_embed.weight"), sizeof(float) * 2 * d, cudaMemcpyDeviceToDevice, st));
  if (ensure_streams() || zero_forward_outputs(st, have_actions_ && !cfg_.no_encoder)) return -1;
  // The CVAE encoder (-> z -> latent token) and the camera backbones (-> image tokens) are independent until the main
  // encoder reads src_: the CVAE branch runs on the side stream so its small launches fill the SMs the convs leave idle.
  cudaStream_t sc = st;
  if (have_actions_ && cfg_.no_encoder) {
    // encoder=None (detr_vae.py:109-112): latent_sample = 0 also in the training call; mu = logvar = None upstream and
    // kl = 0 (policy.py:306-307) -- zero mu / logvar give exactly that in the loss kernel (and zero KL gradients)
    ACT_CHECK(is_pad != nullptr, "training call needs is_pad");
    ACT_CUDA(cudaMemsetAsync(z_, 0, sizeof(float) * B * Z, st));
    ACT_CUDA(cudaMemsetAsync(mu_, 0, sizeof(float) * B * Z, st));
    ACT_CUDA(cudaMemsetAsync(logvar_, 0, sizeof(float) * B * Z, st));
  } else if (have_actions_) {
    ACT_CHECK(is_pad && eps, "training call needs is_pad and eps");
    if (fork_side(st)) return -1;
    if (concurrent_) sc = side_;
    if (cvae_fwd(B, qpos, actions, actions_bs, is_pad, is_pad_bs, eps, sc)) return -1;
  } else {
    ACT_CUDA(cudaMemsetAsync(z_, 0, sizeof(float) * B * Z, st));             // latent_sample = 0 (detr_vae.py:157)
  }
  SgemmArgs g;
  g.I = B; g.J = d; g.R = Z; g.A = z_; g.ai = act::lin(Z); g.ar = act::lin(1);
  g.B = P("latent_out_proj.weight"); g.br = act::lin(1); g.bj = act::lin(Z); g.bias = P("latent_out_proj.bias");
  g.C = src_; g.ci = act::lin(1LL * L_ * d); g.cj = act::lin(1);
  RUN(sgemm_generic(g, sc));
  g = SgemmArgs(); g.I = B; g.J = d; g.R = S; g.A = qpos; g.ai = act::lin(S); g.ar = act::lin(1);
  g.B = P("input_proj_robot_state.weight"); g.br = act::lin(1); g.bj = act::lin(S); g.bias = P("input_proj_robot_state.bias");
  g.C = src_ + d; g.ci = act::lin(1LL * L_ * d); g.cj = act::lin(1);
  if (qnorm_on_ && !have_actions_) { g.a_mean = qnorm_mean_; g.a_std = qnorm_std_; }     // raw qpos from the rollout loop
  RUN(sgemm_generic(g, st));
  if (sc != st) tc_set_sm_reserve(sm_reserve_);
  for (int k = 0; k < cfg_.num_cameras; ++k)
    if (backbone_fwd(k, B, image, st)) { tc_set_sm_reserve(0); return -1; }
  tc_set_sm_reserve(0);
  if (sc != st && join_side(st)) return -1;
  RUN(add_pos_cast(src_, pos_main_, L_, enc_a_[0].xb, enc_a_[0].xpb, B * L_, d, st));
  // the decoder's query self-attention block reads weights only: lane 1 runs it beside the main encoder
  cudaStream_t dl = aux_fork(st, 1);
  if (dl != st && dec_self_fwd(B, dl)) return -1;
  const int nl = cfg_.enc_layers;
  for (int l = 0; l < nl; ++l) {
    const float* x = l == 0 ? src_ : enc_a_[l - 1].x2;
    const bool more = l + 1 < nl;
    if (enc_layer_fwd(enc_w_[l], enc_a_[l], x, pos_main_, pos_main_, L_, nullptr, B, true, more ? enc_a_[l + 1].xb : mem_b_,
                      more ? enc_a_[l + 1].xpb : mem_pb_, st, SITE_ENC + 8 * l)) return -1;
  }
  if (dl == st) { if (dec_self_fwd(B, st)) return -1; }
  else if (aux_join(st, 1)) return -1;
  if (dec_cross_fwd(B, st)) return -1;
  const int Md = B * nq;
  g = SgemmArgs(); g.I = Md; g.J = A; g.R = d; g.A = hs_; g.ai = act::lin(d); g.ar = act::lin(1);
  g.B = P("action_head.weight"); g.br = act::lin(1); g.bj = act::lin(d); g.bias = P("action_head.bias");
  g.C = a_hat_; g.ci = act::lin(A); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  if (is_pad_hat) {
    g = SgemmArgs(); g.I = Md; g.J = 1; g.R = d; g.A = hs_; g.ai = act::lin(d); g.ar = act::lin(1);
    g.B = P("is_pad_head.weight"); g.br = act::lin(1); g.bj = act::lin(d); g.bias = P("is_pad_head.bias");
    g.C = is_pad_hat_; g.ci = act::lin(1); g.cj = act::lin(1);
    RUN(sgemm_generic(g, st));
    ACT_CUDA(cudaMemcpyAsync(is_pad_hat, is_pad_hat_, sizeof(float) * Md, cudaMemcpyDeviceToDevice, st));
  }
  if (have_actions_) {
    RUN(act_loss(a_hat_, actions, actions_bs, is_pad, is_pad_bs, mu_, logvar_, B, nq, A, Z, cfg_.kl_weight, 1.f, losses_,
                 nullptr, nullptr, nullptr, st));
    if (losses) ACT_CUDA(cudaMemcpyAsync(losses, losses_, sizeof(float) * 3, cudaMemcpyDeviceToDevice, st));
    if (mu) ACT_CUDA(cudaMemcpyAsync(mu, mu_, sizeof(float) * B * Z, cudaMemcpyDeviceToDevice, st));
    if (logvar) ACT_CUDA(cudaMemcpyAsync(logvar, logvar_, sizeof(float) * B * Z, cudaMemcpyDeviceToDevice, st));
  }
  if (a_hat) ACT_CUDA(cudaMemcpyAsync(a_hat, a_hat_, sizeof(float) * Md * A, cudaMemcpyDeviceToDevice, st));
  return 0;
}

}  // namespace act
