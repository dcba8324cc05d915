// Engine: backward pass.  Hand-derived reverse of engine_fwd.cu (the reference relies on torch autograd,
// imitate_episodes.py:605-606).  Every weight gradient is accumulated with fp32 atomics into the zeroed flat
// gradient buffer, so split-K, shared weights (input_proj across cameras) and stage-wise execution all compose.
#include "engine.h"
#include <cmath>

namespace act {

#define RUN(expr) do { ++launches_; if ((expr) != 0) return -1; } while (0)

static int roundup(int a, int b) { return (a + b - 1) / b * b; }

__global__ void add_out_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long long n4) {
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * 1LL * blockDim.x) {
    const float4 x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
    reinterpret_cast<float4*>(out)[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
  }
}
static int add_out(const float* a, const float* b, float* out, long long n, cudaStream_t st) {
  const long long n4 = n / 4;
  add_out_kernel<<<static_cast<int>(std::min<long long>((n4 + 255) / 256, 148 * 8)), 256, 0, st>>>(a, b, out, n4);
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// gradient rows of one camera's feature tokens: src fp32 [B, L, d] rows (2 + y*WfK + cam*Wf + x) -> bf16 [B*Hf*Wf, d]
__global__ void gather_tokens_kernel(const float* __restrict__ src, bf16* __restrict__ out, int B, int L, int Hf, int Wf,
                                     int K, int cam, int d4) {
  const long long total = 1LL * B * Hf * Wf * d4;
  for (long long i = blockIdx.x * 1LL * blockDim.x + threadIdx.x; i < total; i += gridDim.x * 1LL * blockDim.x) {
    const int c = static_cast<int>(i % d4);
    long long r = i / d4;
    const int x = static_cast<int>(r % Wf); r /= Wf;
    const int y = static_cast<int>(r % Hf);
    const int b = static_cast<int>(r / Hf);
    const long long row = 1LL * b * L + 2 + (1LL * y * K + cam) * Wf + x;
    const float4 v = reinterpret_cast<const float4*>(src)[row * d4 + c];
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo); o.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(out)[i] = o;
  }
}

// conv1 weight gradient epilogue: dw[co][c][kh][kw] += scale[co] * T[kh][co][kw*4 + c]  (staged layout of tc_wgrad)
__global__ void conv1_grad_fold_kernel(const float* __restrict__ tmp, const float* __restrict__ scale, float* __restrict__ dw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 147) return;
  const int co = i / 147, k = i % 147, c = k / 49, kh = (k % 49) / 7, kw = k % 7;
  dw[i] += scale[co] * tmp[(kh * 64 + co) * 64 + kw * 4 + c];
}

// dX[M,K] = dY[M,N] * W[N,K]   (A = dY K-major, B = W read MN-major)
int Engine::linear_dx(const bf16* dY, int M, const LinW& w, uint32_t flags, const float* res_f32, const bf16* aux,
                      float* out_f32, bf16* out_bf16, cudaStream_t st, bool splitk, float alpha) {
  MatDesc A, B;
  A.ptr = dY; A.ld = w.N; B.ptr = w.w; B.mn_major = 1; B.ld = w.K;
  EpiDesc e;
  e.alpha = alpha;
  e.flags = flags | (out_f32 ? EPI_OUT_F32 : 0) | (out_bf16 ? EPI_OUT_BF16 : 0) | (res_f32 ? EPI_RES_F32 : 0) | (aux ? EPI_MASK_POS : 0);
  e.res_f32 = res_f32; e.out_f32 = out_f32; e.out_bf16 = out_bf16; e.aux_bf16 = aux; e.ldc = w.K; e.ldr = w.K;
  int bn = 0, ks = 1;
  if (splitk && splitk_on() && out_f32 && !out_bf16 && !aux && !(flags & EPI_RELU)) {   // caller pre-zeroed out_f32
    plan_splitk(M, w.K, w.N, &bn, &ks);
    if (ks > 1) e.flags |= EPI_ATOMIC; else bn = 0;
  }
  ++launches_;
  return tc_gemm(M, w.K, w.N, 1, 1, A, B, e, ks, st, bn);
}

// dW[N,K] += dY^T X ; db[N] += colsum(dY)
// shared_weight: the same weight takes gradient from several calls of one step (input_proj, once per camera)
int Engine::linear_dwb(const bf16* dY, int M, const LinW& w, const bf16* X, cudaStream_t st, bool shared_weight) {
  MatDesc A, B;
  A.ptr = dY; A.mn_major = 1; A.ld = w.N; B.ptr = X; B.mn_major = 1; B.ld = w.K;
  EpiDesc e;
  e.flags = EPI_OUT_F32 | EPI_ATOMIC; e.out_f32 = w.gw; e.ldc = w.K;
  int bn = 64, ks = 1;
  plan_splitk(w.N, w.K, M, &bn, &ks);
  // one split and one contributor: plain (sector-wide) stores into the zeroed gradient instead of 6.5 MB of fp32 L2 atomics
  if (ks == 1 && !shared_weight) e.flags = EPI_OUT_F32;
  ++launches_;
  cudaStream_t wg = wg_fork(st);                       // dY (from the arena) is final on st at this point
  if (tc_gemm(w.N, w.K, M, 1, 1, A, B, e, ks, wg, bn)) return -1;
  if (w.gb) RUN(colsum_bf16(dY, w.N, M, w.N, w.gb, wg));
  return 0;
}

int Engine::attn_bwd(const bf16* dO, const bf16* Q, long long ldq, const bf16* K, long long ldk, const bf16* V, long long ldv,
                     const bf16* P, bf16* dQ, long long lddq, bf16* dK, long long lddk, bf16* dV, long long lddv, int B,
                     int Lq, int Lk, cudaStream_t st, float* S, bf16* dS, bool use_aux, uint32_t drop, const bf16* P2, const bf16* O) {
  if (!S) S = Sbuf_;
  if (!dS) dS = dSbuf_;
  const bool dropping = drop && drop_active_;      // O = dropout(P) V: dV sees the dropped weights, dP is masked again
  ACT_CHECK(!dropping || P2, "attention dropout backward needs the dropped probabilities");
  const int h = cfg_.nheads, hd = 64, Lkp = roundup(Lk, 8);
  const long long ph = 1LL * Lq * Lkp, pb = 1LL * h * Lq * Lkp;
  // dV and dK only feed the projections' gradients: they run on the aux lane beside dP -> softmax backward -> dQ
  cudaStream_t ax = use_aux ? aux_fork(st) : st;
  {  // dV[j,d] = sum_i P[i,j] dO[i,d]
    MatDesc a, b; EpiDesc e;
    a.ptr = dropping ? P2 : P; a.mn_major = 1; a.ld = Lkp; a.sz0 = ph; a.sz1 = pb;
    b.ptr = dO; b.mn_major = 1; b.ld = 512; b.sz0 = hd; b.sz1 = 1LL * Lq * 512;
    e.flags = EPI_OUT_BF16; e.out_bf16 = dV; e.ldc = lddv; e.osz0 = hd; e.osz1 = Lk * lddv;
    ++launches_;
    if (tc_gemm(Lk, hd, Lq, h, B, a, b, e, 1, ax)) return -1;
  }
  {  // dP[i,j] = sum_d dO[i,d] V[j,d], then dS = scale * P o (dP - rowsum(dP o P))
    MatDesc a, b; EpiDesc e;
    a.ptr = dO; a.ld = 512; a.sz0 = hd; a.sz1 = 1LL * Lq * 512;
    b.ptr = V; b.ld = ldv; b.sz0 = hd; b.sz1 = Lk * ldv;
    if (tc_attn_scores_applicable(Lk, Lkp, dropping)) {
      ++launches_;
      const DropDesc dd = dropping ? drop_site(drop) : DropDesc();
      MatDesc od; od.ptr = O; od.ld = 512; od.sz0 = hd; od.sz1 = 1LL * Lq * 512;      // rowsum(P o dP) = dO . O (long rows: one pass)
      if (tc_attn_scores(1, a, b, Lq, Lk, Lkp, h, B, 0.125f, nullptr, P, dS, st, dropping ? &dd : nullptr, nullptr, O ? &od : nullptr)) return -1;
    } else {
      ACT_CHECK(!dropping, "attention dropout is implemented in the fused scores kernels only (ACT_FUSED_ATTN=0 is set)");
      e.flags = EPI_OUT_F32; e.out_f32 = S; e.ldc = Lkp; e.osz0 = ph; e.osz1 = pb;
      ++launches_;
      if (tc_gemm(Lq, Lk, hd, h, B, a, b, e, 1, st)) return -1;
      RUN(softmax_bwd(P, S, dS, B * h * Lq, Lk, Lkp, 0.125f, st));
    }
  }
  if (ax != st) ax = aux_fork(st);                   // dS is final
  {  // dK[j,d] = sum_i dS[i,j] Q[i,d]
    MatDesc a, b; EpiDesc e;
    a.ptr = dS; a.mn_major = 1; a.ld = Lkp; a.sz0 = ph; a.sz1 = pb;
    b.ptr = Q; b.mn_major = 1; b.ld = ldq; b.sz0 = hd; b.sz1 = Lq * ldq;
    e.flags = EPI_OUT_BF16; e.out_bf16 = dK; e.ldc = lddk; e.osz0 = hd; e.osz1 = Lk * lddk;
    ++launches_;
    if (tc_gemm(Lk, hd, Lq, h, B, a, b, e, 1, ax)) return -1;
  }
  {  // dQ[i,d] = sum_j dS[i,j] K[j,d]
    MatDesc a, b; EpiDesc e;
    a.ptr = dS; a.ld = Lkp; a.sz0 = ph; a.sz1 = pb;
    b.ptr = K; b.mn_major = 1; b.ld = ldk; b.sz0 = hd; b.sz1 = Lk * ldk;
    e.flags = EPI_OUT_BF16; e.out_bf16 = dQ; e.ldc = lddq; e.osz0 = hd; e.osz1 = Lq * lddq;
    ++launches_;
    if (tc_gemm(Lq, hd, Lk, h, B, a, b, e, 1, st)) return -1;
  }
  if (ax != st && aux_join(st)) return -1;
  return 0;
}

// Reverse of enc_layer_fwd.  Input gradient = dx2 (+ dx2b); outputs dx_out (wrt x) and gxp_out (wrt x + pos, i.e. the
// q/k branch) -- the caller adds them (and reduces gxp_out into learned position embeddings where they exist).
int Engine::enc_layer_bwd(const EncW& w, EncActs& a, const float* x_in, const float* dx2, const float* dx2b, float* dx_out,
                          float* gxp_out, int L, int B, cudaStream_t st, uint32_t site) {
  (void)x_in;
  const int M = B * L, d = cfg_.hidden_dim, ff = cfg_.dim_ff;
  // train-mode dropout: the gradient handed to a linear layer whose output was dropped is masked again (regenerated
  // mask); the saved FFN hidden activation already carries its mask (zeros), so its backward only needs the 1/(1-p)
  const DropDesc d1 = drop_site(site + 1), d3 = drop_site(site + 3);
  const DropDesc* pd1 = drop_active_ ? &d1 : nullptr;
  const DropDesc* pd3 = drop_active_ ? &d3 : nullptr;
  const float hs = drop_active_ ? 1.f / (1.f - drop_p_) : 1.f;
  bf16 *g1 = gy_alloc(1LL * M * d), *ghh = gy_alloc(1LL * M * ff), *g1b = gy_alloc(1LL * M * d);
  bf16 *gqk = gy_alloc(2LL * M * d), *gv = gy_alloc(1LL * M * d);
  if (!g1 || !ghh || !g1b || !gqk || !gv) return -1;
  RUN(ln_bwd(dx2, dx2b, a.t2, a.mean2, a.rstd2, w.n2.g, gA_, g1, w.n2.gg, w.n2.gb, M, d, st, pd3));
  if (linear_dwb(g1, M, w.l2, a.hh, st)) return -1;
  if (linear_dx(g1, M, w.l2, 0, nullptr, a.hh, nullptr, ghh, st, false, hs)) return -1;
  if (linear_dwb(ghh, M, w.l1, a.x1b, st)) return -1;
  float* gBz = splitk_on() ? gz_alloc(1LL * M * d) : gB_;        // zeroed slice: the K = 3200 data gradient runs split-K
  if (!gBz) return -1;
  if (zero_b_pending_) ACT_CUDA(cudaStreamWaitEvent(st, ev_zero_b_, 0));
  if (linear_dx(ghh, M, w.l1, 0, gA_, nullptr, gBz, nullptr, st, true)) return -1;
  RUN(ln_bwd(gBz, nullptr, a.t1, a.mean1, a.rstd1, w.n1.g, gA_, g1b, w.n1.gg, w.n1.gb, M, d, st, pd1));
  if (linear_dwb(g1b, M, w.sa.o, a.o, st)) return -1;
  if (linear_dx(g1b, M, w.sa.o, 0, nullptr, nullptr, nullptr, gb2_, st)) return -1;
  if (attn_bwd(gb2_, a.qk, 2 * d, a.qk + d, 2 * d, a.v, d, a.P, gqk, 2 * d, gqk + d, 2 * d, gv, d, B, L, L, st, nullptr, nullptr,
               true, site + 0, a.P2, a.o)) return -1;
  if (linear_dwb(gqk, M, w.sa.qk, a.xpb, st)) return -1;
  if (linear_dwb(gv, M, w.sa.v, a.xb, st)) return -1;
  cudaStream_t ax = aux_fork(st);                    // the two input gradients are independent
  if (linear_dx(gv, M, w.sa.v, 0, gA_, nullptr, dx_out, nullptr, ax)) return -1;
  if (linear_dx(gqk, M, w.sa.qk, 0, nullptr, nullptr, gxp_out, nullptr, st)) return -1;
  if (ax != st && aux_join(st)) return -1;
  return 0;
}


// Heads, final norm, decoder layer 0, main encoder, then the gradients of the token-producing projections.
int Engine::transformer_bwd(float loss_grad, cudaStream_t st) {
  const int B = cur_B_, d = cfg_.hidden_dim, nq = cfg_.num_queries, A = cfg_.action_dim, Z = cfg_.latent_dim;
  const int Md = B * nq, Me = B * L_;
  const DecW& w = dec_w_[0];
  DecActs& a = dec_a_;
  RUN(act_loss(a_hat_, in_actions_, in_actions_bs_, in_is_pad_, in_is_pad_bs_, mu_, logvar_, B, nq, A, Z, cfg_.kl_weight,
               loss_grad, losses_ + 4, d_a_hat_, d_mu_, d_logvar_, st, loss_gvec_));
  SgemmArgs g;
  // action_head: dW[a,c] += sum_m d_a[m,a] hs[m,c]; db; d_hs = d_a W
  g = SgemmArgs(); g.I = A; g.J = d; g.R = Md; g.A = d_a_hat_; g.ai = act::lin(1); g.ar = act::lin(A);
  g.B = hs_; g.br = act::lin(d); g.bj = act::lin(1); g.C = G("action_head.weight"); g.ci = act::lin(d); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, st));
  g = SgemmArgs(); g.I = 1; g.J = A; g.R = Md; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = d_a_hat_; g.br = act::lin(A); g.bj = act::lin(1); g.C = G("action_head.bias"); g.ci = act::lin(0); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, st));
  g = SgemmArgs(); g.I = Md; g.J = d; g.R = A; g.A = d_a_hat_; g.ai = act::lin(A); g.ar = act::lin(1);
  g.B = P("action_head.weight"); g.br = act::lin(d); g.bj = act::lin(1); g.C = d_hs_; g.ci = act::lin(d); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  // shared final norm, norm3, FFN  (every dY that a deferred weight gradient reads lives in the arena)
  gy_off_ = 0; gy_lim_ = gy_cvae_base_;
  gz_off_ = 0;
  const int ff = cfg_.dim_ff;
  bf16 *g3 = gy_alloc(1LL * Md * d), *ghh = gy_alloc(1LL * Md * ff), *g2 = gy_alloc(1LL * Md * d), *gq = gy_alloc(1LL * Md * d);
  bf16 *gk = gy_alloc(1LL * Me * d), *gcv = gy_alloc(1LL * Me * d), *g1 = gy_alloc(1LL * Md * d), *gqk = gy_alloc(2LL * Md * d);
  bf16 *gv = gy_alloc(1LL * Md * d);
  if (!g3 || !ghh || !g2 || !gq || !gk || !gcv || !g1 || !gqk || !gv) return -1;
  RUN(ln_bwd(d_hs_, nullptr, a.y3, hs_mean_, hs_rstd_, dec_norm_.g, gA_, nullptr, dec_norm_.gg, dec_norm_.gb, Md, d, st));
  const DropDesc dd1 = drop_site(SITE_DEC + 1), dd3 = drop_site(SITE_DEC + 3), dd5 = drop_site(SITE_DEC + 5);
  const DropDesc *pdd1 = drop_active_ ? &dd1 : nullptr, *pdd3 = drop_active_ ? &dd3 : nullptr, *pdd5 = drop_active_ ? &dd5 : nullptr;
  const float hs = drop_active_ ? 1.f / (1.f - drop_p_) : 1.f;
  RUN(ln_bwd(gA_, nullptr, a.t3, a.mean3, a.rstd3, w.n3.g, gB_, g3, w.n3.gg, w.n3.gb, Md, d, st, pdd3));
  if (linear_dwb(g3, Md, w.l2, a.hh, st)) return -1;
  if (linear_dx(g3, Md, w.l2, 0, nullptr, a.hh, nullptr, ghh, st, false, hs)) return -1;
  if (linear_dwb(ghh, Md, w.l1, a.y2b, st)) return -1;
  float* gAz = splitk_on() ? gz_alloc(1LL * Md * d) : gA_;
  if (!gAz) return -1;
  if (zero_b_pending_) ACT_CUDA(cudaStreamWaitEvent(st, ev_zero_b_, 0));
  if (linear_dx(ghh, Md, w.l1, 0, gB_, nullptr, gAz, nullptr, st, true)) return -1;
  // norm2, cross-attention.  dgB_ (wrt y1) and dgE_ (wrt y1 + query_pos) are private to the decoder: the query
  // self-attention backward below consumes them on lane 1 while this stream goes on with the encoder layers.
  RUN(ln_bwd(gAz, nullptr, a.t2, a.mean2, a.rstd2, w.n2.g, dgB_, g2, w.n2.gg, w.n2.gb, Md, d, st, pdd5));
  if (linear_dwb(g2, Md, w.ca.o, a.co, st)) return -1;
  if (linear_dx(g2, Md, w.ca.o, 0, nullptr, nullptr, nullptr, gb2_, st)) return -1;
  if (attn_bwd(gb2_, a.q, d, a.k, d, a.cv, d, a.cP, gq, d, gk, d, gcv, d, B, nq, L_, st, nullptr, nullptr, true, SITE_DEC + 4, a.cP2, a.co)) return -1;
  if (linear_dwb(gq, Md, w.ca.q, a.y1pb, st)) return -1;
  if (linear_dx(gq, Md, w.ca.q, 0, nullptr, nullptr, dgE_, nullptr, st)) return -1;          // wrt y1 + query_pos
  if (linear_dwb(gk, Me, w.ca.k, mem_pb_, st)) return -1;
  if (linear_dwb(gcv, Me, w.ca.v, mem_b_, st)) return -1;
  float* g_addpos = G("additional_pos_embed.weight");
  float* g_query = G("query_embed.weight");
  auto reduce_batch_on = [&](cudaStream_t s2, const float* src, long long batch_stride, int n, float* dst) {   // dst[j] += sum_b src[b*bs + j]
    SgemmArgs r;
    r.I = 1; r.J = n; r.R = B; r.A = one_; r.ai = act::lin(0); r.ar = act::lin(0);
    r.B = src; r.br = act::lin(batch_stride); r.bj = act::lin(1); r.C = dst; r.ci = act::lin(0); r.cj = act::lin(1); r.accumulate = 1;
    ++launches_;
    return sgemm_generic(r, s2);
  };
  auto reduce_batch = [&](const float* src, long long batch_stride, int n, float* dst) { return reduce_batch_on(st, src, batch_stride, n, dst); };
  {   // query self-attention block: a leaf of the backward graph (its only consumers are weight / query_embed gradients)
    cudaStream_t dl = aux_fork(st, 1);
    if (reduce_batch_on(dl, dgE_, 1LL * nq * d, nq * d, g_query)) return -1;
    RUN(ln_bwd(dgB_, dgE_, a.t1, a.mean1, a.rstd1, w.n1.g, dgA_, g1, w.n1.gg, w.n1.gb, Md, d, dl, pdd1));
    if (linear_dwb(g1, Md, w.sa.o, a.o, dl)) return -1;
    if (linear_dx(g1, Md, w.sa.o, 0, nullptr, nullptr, nullptr, dgb2_, dl)) return -1;
    if (attn_bwd(dgb2_, a.qk, 2 * d, a.qk + d, 2 * d, a.v, d, a.P, gqk, 2 * d, gqk + d, 2 * d, gv, d, B, nq, nq, dl, Sbuf2_, dSbuf2_, false,
                 SITE_DEC + 0, a.P2, a.o)) return -1;
    if (linear_dwb(gqk, Md, w.sa.qk, a.xpb, dl)) return -1;
    if (linear_dwb(gv, Md, w.sa.v, a.xb, dl)) return -1;
    if (linear_dx(gqk, Md, w.sa.qk, 0, nullptr, nullptr, dgE_, nullptr, dl)) return -1;
    if (reduce_batch_on(dl, dgE_, 1LL * nq * d, nq * d, g_query)) return -1;
  }
  {   // memory gradients: the two cross-attention input projections are independent
    cudaStream_t ax = aux_fork(st);
    if (linear_dx(gcv, Me, w.ca.v, 0, nullptr, nullptr, gMem_, nullptr, ax)) return -1;         // wrt memory
    if (linear_dx(gk, Me, w.ca.k, 0, nullptr, nullptr, gMemPos_, nullptr, st)) return -1;       // wrt memory + pos
    if (ax != st && aux_join(st)) return -1;
  }
  if (reduce_batch(gMemPos_, 1LL * L_ * d, 2 * d, g_addpos)) return -1;
  // main encoder, last layer first
  const float* dx2 = gMem_;
  const float* dx2b = gMemPos_;
  for (int l = cfg_.enc_layers - 1; l >= 0; --l) {
    if (enc_layer_bwd(enc_w_[l], enc_a_[l], nullptr, dx2, dx2b, gC_, gD_, L_, B, st, SITE_ENC + 8 * l)) return -1;
    if (reduce_batch(gD_, 1LL * L_ * d, 2 * d, g_addpos)) return -1;
    dx2 = gC_; dx2b = gD_;
  }
  RUN(add_out(gC_, gD_, d_src_, 1LL * Me * d, st));
  // token 0 <- latent_out_proj(z), token 1 <- input_proj_robot_state(qpos)
  const long long ts = 1LL * L_ * d;
  g = SgemmArgs(); g.I = d; g.J = Z; g.R = B; g.A = d_src_; g.ai = act::lin(1); g.ar = act::lin(ts);
  g.B = z_; g.br = act::lin(Z); g.bj = act::lin(1); g.C = G("latent_out_proj.weight"); g.ci = act::lin(Z); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, st));
  if (reduce_batch(d_src_, ts, d, G("latent_out_proj.bias"))) return -1;
  g = SgemmArgs(); g.I = B; g.J = Z; g.R = d; g.A = d_src_; g.ai = act::lin(ts); g.ar = act::lin(1);
  g.B = P("latent_out_proj.weight"); g.br = act::lin(Z); g.bj = act::lin(1); g.C = dz_; g.ci = act::lin(Z); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  const int S = cfg_.state_dim;
  g = SgemmArgs(); g.I = d; g.J = S; g.R = B; g.A = d_src_ + d; g.ai = act::lin(1); g.ar = act::lin(ts);
  g.B = in_qpos_; g.br = act::lin(S); g.bj = act::lin(1); g.C = G("input_proj_robot_state.weight"); g.ci = act::lin(S); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, st));
  if (reduce_batch(d_src_ + d, ts, d, G("input_proj_robot_state.bias"))) return -1;
  // image tokens <- input_proj(layer4 features) per camera
  const int Mf = B * Hf_ * Wf_;
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    const long long total = 1LL * Mf * (d / 4);
    bf16* gtok = gy_alloc(1LL * Mf * d);
    if (!gtok) return -1;
    gather_tokens_kernel<<<static_cast<int>(std::min<long long>((total + 255) / 256, 148 * 8)), 256, 0, st>>>(
        d_src_, gtok, B, L_, Hf_, Wf_, cfg_.num_cameras, k, d / 4);
    ++launches_;
    ACT_CUDA(cudaGetLastError());
    const bf16* feat = cam_a_[k].blk[7].y;
    if (linear_dwb(gtok, Mf, input_proj_, feat, st, cfg_.num_cameras > 1)) return -1;
    if (linear_dx(gtok, Mf, input_proj_, 0, nullptr, feat, nullptr, d_feat_ + 1LL * k * cfg_.batch_max * Hf_ * Wf_ * 512, st)) return -1;
  }
  return aux_join(st, 1);                            // the decoder self-attention backward (lane 1)
}

// Reverse of cvae_fwd: reparametrisation + KL, latent_proj, the 4 masked encoder layers, the token projections.
int Engine::cvae_bwd(cudaStream_t st) {
  const int B = cur_B_, d = cfg_.hidden_dim, nq = cfg_.num_queries, A = cfg_.action_dim, Z = cfg_.latent_dim, S = cfg_.state_dim;
  const int Mc = B * Lc_, nl = cfg_.enc_layers;
  const long long ts = 1LL * Lc_ * d;
  gy_off_ = gy_cvae_base_; gy_lim_ = gy_cap_;
  gz_off_ = gz_cvae_base_;
  RUN(reparam_bwd(dz_, eps_, logvar_, d_mu_, d_logvar_, d_info_, B, Z, st));
  SgemmArgs g;
  const float* cls_out = cvae_a_[nl - 1].x2;
  g = SgemmArgs(); g.I = 2 * Z; g.J = d; g.R = B; g.A = d_info_; g.ai = act::lin(1); g.ar = act::lin(2 * Z);
  g.B = cls_out; g.br = act::lin(ts); g.bj = act::lin(1); g.C = G("latent_proj.weight"); g.ci = act::lin(d); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));                      // parameter gradients only: beside the chain, on the weight-gradient lanes
  g = SgemmArgs(); g.I = 1; g.J = 2 * Z; g.R = B; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = d_info_; g.br = act::lin(2 * Z); g.bj = act::lin(1); g.C = G("latent_proj.bias"); g.ci = act::lin(0); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  // gradient wrt the encoder output: only the cls rows are non-zero
  ACT_CUDA(cudaMemsetAsync(gC_, 0, sizeof(float) * Mc * d, st));
  g = SgemmArgs(); g.I = B; g.J = d; g.R = 2 * Z; g.A = d_info_; g.ai = act::lin(2 * Z); g.ar = act::lin(1);
  g.B = P("latent_proj.weight"); g.br = act::lin(d); g.bj = act::lin(1); g.C = gC_; g.ci = act::lin(ts); g.cj = act::lin(1);
  RUN(sgemm_generic(g, st));
  const float* dx2 = gC_;
  const float* dx2b = nullptr;
  for (int l = nl - 1; l >= 0; --l) {
    // data-parallel step: collectives run in enqueue order, so the backbone branch (other stream) is enqueued at the
    // point of this branch where its gradients become final -- after dp_bb_slot_ encoder layers
    if (dp_bb_pending_ && nl - 1 - l == dp_bb_slot_ && backbone_bwd_and_step(dp_main_st_)) return -1;
    if (enc_layer_bwd(cvae_w_[l], cvae_a_[l], nullptr, dx2, dx2b, gC_, gD_, Lc_, B, st, SITE_CVAE + 8 * l)) return -1;
    dx2 = gC_; dx2b = gD_;
    if (inline_opt_on_ && l < 16 && l >= dp_inline_from_ && cvae_layer_n_[l] > 0) {
      // this layer's gradients are final and nothing below reads its weights: update it under the remaining layers
      cudaStream_t wg = wg_gather(st, NWG);
      if (grads_ready(cvae_layer_off_[l], cvae_layer_n_[l], wg)) return -1;
      if (adamw_range(cvae_layer_off_[l], cvae_layer_n_[l], cfg_.lr, true, inline_opt_step_, wg, 3)) return -1;
    }
  }
  RUN(add_out(gC_, gD_, d_tokC_, 1LL * Mc * d, st));
  // token 0: cls_embed ; token 1: encoder_joint_proj(qpos) ; tokens 2..: encoder_action_proj(actions).  Five independent tiny
  // reductions at the very end of the branch: spread over the weight-gradient lanes (the optimiser lane gathers them)
  g = SgemmArgs(); g.I = 1; g.J = d; g.R = B; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = d_tokC_; g.br = act::lin(ts); g.bj = act::lin(1); g.C = G("cls_embed.weight"); g.ci = act::lin(0); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  g = SgemmArgs(); g.I = d; g.J = S; g.R = B; g.A = d_tokC_ + d; g.ai = act::lin(1); g.ar = act::lin(ts);
  g.B = in_qpos_; g.br = act::lin(S); g.bj = act::lin(1); g.C = G("encoder_joint_proj.weight"); g.ci = act::lin(S); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  g = SgemmArgs(); g.I = 1; g.J = d; g.R = B; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = d_tokC_ + d; g.br = act::lin(ts); g.bj = act::lin(1); g.C = G("encoder_joint_proj.bias"); g.ci = act::lin(0); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  // dW[c][a] += sum_t dTok[t][c] * actions[t][a], issued transposed (rows = a, columns = c) so lanes run along c
  g = SgemmArgs(); g.I = A; g.J = d; g.R = B * nq; g.A = in_actions_; g.ai = act::lin(1); g.ar = batched(nq, in_actions_bs_, A);
  g.B = d_tokC_ + 2 * d; g.br = batched(nq, ts, d); g.bj = act::lin(1); g.C = G("encoder_action_proj.weight"); g.ci = act::lin(1); g.cj = act::lin(A); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  g = SgemmArgs(); g.I = 1; g.J = d; g.R = B * nq; g.A = one_; g.ai = act::lin(0); g.ar = act::lin(0);
  g.B = d_tokC_ + 2 * d; g.br = batched(nq, ts, d); g.bj = act::lin(1); g.C = G("encoder_action_proj.bias"); g.ci = act::lin(0); g.cj = act::lin(1); g.accumulate = 1;
  RUN(sgemm_generic(g, wg_fork(st)));
  return 0;
}


// Reverse of backbone_fwd for one camera.  The incoming gradient (d_feat_) is already masked by the ReLU of layer4.
// FrozenBN: wrt the conv output the gradient is scale[co] * g; the scale is folded into the packed dgrad weights
// (wpkt) and applied per output channel in the wgrad epilogue, so no elementwise pass is needed.
int Engine::backbone_bwd(int cam, int B, cudaStream_t st) {
  CamW& cw = cam_w_[cam];
  CamActs& ca = cam_a_[cam];
  const bf16* g = d_feat_ + 1LL * cam * cfg_.batch_max * Hf_ * Wf_ * 512;
  // conv weight gradients accumulate in staged layout (line-coalesced reductions) and are folded into dW by ONE launch;
  // all of it runs on the weight-gradient stream, reading activation gradients that live in the conv-gradient arena
  cudaStream_t wg = wg_fork(st, 0);
  ACT_CUDA(cudaMemsetAsync(wstage_, 0, sizeof(float) * wstage_n_, wg));
  if (wg != st) ACT_CUDA(cudaEventRecord(ev_stage_, wg));
  FoldJobs folds;
  long long woff = 0;
  auto wgrad = [&](const ConvShape& cs, const bf16* xin, const bf16* dy, const ConvW& cw2) -> int {
    const long long n = 1LL * cs.KH * cs.KH * cs.Cin * cs.Cout;
    ACT_CHECK(woff + n <= wstage_n_ && folds.n < FoldJobs::MAX, "weight-gradient staging overflow");
    float* T = wstage_ + woff;
    woff += n;
    folds.job[folds.n++] = FoldJob{T, cw2.scale, cw2.gw, 1.f, cs.Cout, cs.Cin, cs.KH * cs.KH, 0};
    ++launches_;
    wg = wg_fork(st);                                // dy is final on st here
    if (wg != st) ACT_CUDA(cudaStreamWaitEvent(wg, ev_stage_, 0));   // staging buffer cleared
    return conv_wgrad_staged(cs, xin, dy, T, wg);
  };
  for (int i = 7; i >= 0; --i) {
    BlockW& bw = cw.blk[i];
    BlockActs& ba = ca.blk[i];
    const int s = i / 2;
    const int Ho = fh_[2 + s], Wo = fw_[2 + s];
    const int st1 = bw.c1.stride;
    const int Hc = (st1 == 2) ? fh_[1 + s] : Ho, Wc = (st1 == 2) ? fw_[1 + s] : Wo;
    const bf16* x_in = (i == 0) ? ca.p1 : ca.blk[i - 1].y;
    bf16* ga = cgy_alloc(1LL * B * Ho * Wo * bw.c1.Cout);
    bf16* dx = cgy_alloc(1LL * B * Hc * Wc * bw.c1.Cin);
    if (!ga || !dx) return -1;
    ConvShape s2{B, Ho, Wo, bw.c2.Cin, bw.c2.Cout, 3, 1, 1};
    ConvShape s1{B, Hc, Wc, bw.c1.Cin, bw.c1.Cout, 3, st1, 1};
    if (wgrad(s2, ba.a, g, bw.c2)) return -1;
    {
      EpiDesc e; e.flags = EPI_OUT_BF16 | EPI_MASK_POS; e.out_bf16 = ga; e.aux_bf16 = ba.a;
      RUN(conv_dgrad(s2, g, bw.c2.wpkt, e, nullptr, st, bw.c2.wpkt2));
    }
    if (wgrad(s1, x_in, ga, bw.c1)) return -1;
    EpiDesc e1; e1.flags = EPI_OUT_BF16; e1.out_bf16 = dx;
    if (i > 0) { e1.flags |= EPI_MASK_POS; e1.aux_bf16 = x_in; }     // x_in = relu output of the previous block
    if (bw.has_ds) {
      ConvShape sd{B, Hc, Wc, bw.ds.Cin, bw.ds.Cout, 1, st1, 0};
      if (wgrad(sd, x_in, g, bw.ds)) return -1;
      ConvShape sdc{B, Ho, Wo, bw.ds.Cin, bw.ds.Cout, 1, 1, 0};      // 1x1 data gradient on the compact (strided) grid
      EpiDesc ec; ec.flags = EPI_OUT_BF16; ec.out_bf16 = cg_[0];
      RUN(conv_dgrad(sdc, g, bw.ds.wpkt, ec, nullptr, st));
      RUN(conv_dgrad(s1, ga, bw.c1.wpkt, e1, cg_[0], st));
      launches_ += 3;                                                // four parity-class launches
    } else {
      e1.flags |= EPI_RES_BF16; e1.res_bf16 = g;
      RUN(conv_dgrad(s1, ga, bw.c1.wpkt, e1, nullptr, st, bw.c1.wpkt2));
    }
    g = dx;
    if (i == 6 && bb_early_ && folds.n > 0) {
      // layer4's weight gradients are all staged: fold them, then exchange + update that range on the optimiser lane
      // while the remaining three stages of the backward run (its packed operands are rebuilt with the others at the end)
      cudaStream_t ol = wg_gather(st, NWG);
      RUN(wgrad_fold_batched(folds, ol));
      folds.n = 0;
      if (grads_ready(bb_l4_off_[cam], bb_l4_n_[cam], ol)) return -1;
      if (adamw_range(bb_l4_off_[cam], bb_l4_n_[cam], cfg_.lr_backbone, false, opt_step_, ol, 3)) return -1;
    }
  }
  wg = wg_gather(st);
  RUN(wgrad_fold_batched(folds, wg));
  if (bb_split_) {
    // single-GPU overlapped step: every weight of this camera except conv1 is final once the fold has run, so its AdamW
    // and the re-pack of its conv operands start here, under the max-pool backward and the stem's weight gradient
    // (conv1 follows in backbone_bwd_and_step); the step's tail loses ~100 us of serial optimiser + re-pack time
    const long long c1 = cw.stem.gw - grads_, c1n = (64LL * 3 * 7 * 7 + 7) / 8 * 8;
    const long long lo = bb_rest_off_[cam], hi = bb_l4_off_[cam] + bb_l4_n_[cam];
    ACT_CHECK(c1 == lo && lo + c1n <= hi, "backbone span does not start with conv1.weight");
    cudaStream_t ol = wg_gather(st, NWG);
    if (adamw_range(lo + c1n, hi - lo - c1n, cfg_.lr_backbone, false, opt_step_, ol, 3)) return -1;
    if (repack_conv_operands(ol, cam, false)) return -1;
  }
  // stem: max-pool backward (+ReLU mask), then dW1 = scale * dy1^T (row windows of the normalised image)
  const int H1 = fh_[1], W1 = fw_[1];
  bf16* gpool = cgy_alloc(1LL * B * H1 * W1 * 64);
  if (!gpool) return -1;
  RUN(maxpool_bwd(g, ca.pidx, nullptr, gpool, B, H1, W1, 64, st));      // x == NULL: the arg-max record carries the ReLU mask
  wg = wg_fork(st);
  // dW1 over the same row-window operand the forward read (ca.xp is still resident): tmp[kh][co][kw*4+c]
  ACT_CUDA(cudaMemsetAsync(conv1_tmp_, 0, sizeof(float) * 64 * 64 * 7, wg));
  {
    WgradLaunch w;
    if (stem_launch(cam, B, nullptr, &w)) return -1;
    w.dy.ptr = gpool; w.dy.C = 64; w.dy.W = W1; w.dy.H = H1; w.dy.N = B;
    w.dy.sW = 64; w.dy.sH = 64LL * W1; w.dy.sN = 64LL * W1 * H1;
    w.dw = conv1_tmp_; w.alpha = 1.f; w.scale = nullptr;
    RUN(tc_wgrad(w, wg));
  }
  conv1_grad_fold_kernel<<<(64 * 147 + 255) / 256, 256, 0, wg>>>(conv1_tmp_, cw.stem.scale, cw.stem.gw);
  ++launches_;
  ACT_CUDA(cudaGetLastError());
  return 0;
}

int Engine::grads_ready(long long off, long long n, cudaStream_t st) {
  if (dp_fused_) return n > 0 ? dp_sync_point(st) : 0;      // every rank's gradients of the range are final past this point
  if (!grads_cb_ || n <= 0) return 0;
  if (grads_cb_(grads_cb_user_, off, n, st)) { set_error("gradient all-reduce callback failed"); return -1; }
  return 0;
}

// Backbone backward of every camera on `st`, then (inside backward_and_step) that region's averaging hook and AdamW on
// the optimiser lane: it trails the batched fold and the stem weight gradient, not the (longer) CVAE branch.
int Engine::backbone_bwd_and_step(cudaStream_t st) {
  dp_bb_pending_ = false;
  cgy_off_ = 0;
  { static int early = -1; if (early < 0) { const char* e = getenv("ACT_BB_EARLY"); early = (e && e[0] == '1') ? 1 : 0; }   // opt-in: measured neutral at N=1, -0.4 % at N=2 (one more barrier site)
    bb_early_ = early != 0 && cfg_.num_cameras <= 8 && concurrent_; }
  { static int split = -1; if (split < 0) { const char* e = getenv("ACT_BB_SPLIT"); split = (e && e[0] == '0') ? 0 : 1; }
    bb_split_ = split != 0 && !bb_early_ && !dp_fused_ && !grads_cb_ && concurrent_ && cfg_.num_cameras <= 8; }
  tc_set_sm_reserve(sm_reserve_);
  for (int k = 0; k < cfg_.num_cameras; ++k)
    if (backbone_bwd(k, cur_B_, st)) { tc_set_sm_reserve(0); bb_split_ = false; return -1; }
  tc_set_sm_reserve(0);
  cudaStream_t wg = wg_gather(st, NWG);
  if (bb_split_) {                                          // what is left: conv1 of every camera and its packed operand
    bb_split_ = false;
    const long long c1n = (64LL * 3 * 7 * 7 + 7) / 8 * 8;
    for (int k = 0; k < cfg_.num_cameras; ++k) {
      if (adamw_range(bb_rest_off_[k], c1n, cfg_.lr_backbone, false, opt_step_, wg, 3)) return -1;
      if (pack_stem_weights(cam_w_[k].stem.w32, cam_w_[k].stem.wpk, wg)) return -1;
      ++launches_;
    }
    return 0;
  }
  if (!bb_early_) {
    if (dp_defer_) {                                        // peer-memory step: one barrier for this and the CVAE head (p2p_finish)
      dp_defer_now_ = true;
      const int rc = adamw_region(R_BACKBONE, opt_step_, wg, 8);
      dp_defer_now_ = false;
      return rc;
    }
    if (grads_ready(region_off_[R_BACKBONE], region_n_[R_BACKBONE], wg)) return -1;
    return adamw_region(R_BACKBONE, opt_step_, wg, dp_fused_ ? 6 : 3);
  }
  bb_early_ = false;
  // the remainder of every camera's span (stem .. layer3): one barrier covers all of them on the peer-memory path
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    if ((!dp_fused_ || k == 0) && grads_ready(bb_rest_off_[k], bb_rest_n_[k], wg)) return -1;
    if (adamw_range(bb_rest_off_[k], bb_rest_n_[k], cfg_.lr_backbone, false, opt_step_, wg, dp_fused_ ? 6 : 3)) return -1;
  }
  if (dp_fused_) { dp_pack_pending_ = true; return 0; }     // re-pack after the closing barrier (every rank's slices landed)
  return repack_conv_operands(wg);
}

int Engine::backward_and_step(int step, float loss_grad, cudaStream_t st, GradsReadyFn cb, void* cb_user, bool p2p) {
  ACT_CHECK(cfg_.train && grads_ != nullptr, "engine has no training workspace / gradient buffer");
  ACT_CHECK(have_actions_ && cur_B_ > 0, "backward needs a preceding training forward (with actions)");
  ACT_CHECK(!p2p || (peers_.world > 1 && dp_stream_ != nullptr), "peer-memory data-parallel step without connected peers");
  ACT_CHECK(!(p2p && cb), "peer-memory data-parallel step and an all-reduce callback are exclusive");
  if (ensure_streams()) return -1;
  if (begin_opt_step(step, st)) return -1;                  // first launch of the step: every optimiser lane forks after it
  // the peer-memory step ends with a barrier: every rank has finished reading this rank's gradients and writing its
  // parameters before the caller's stream goes on (next zero_grad / forward)
  auto p2p_finish = [&]() -> int {
    dp_defer_ = dp_defer_now_ = false;
    if (dp_flush_deferred()) return -1;                    // backbone + CVAE head / tail behind one barrier
    if (dp_sync_point(st)) return -1;                      // (also starts a still-pending conv re-pack)
    ACT_CUDA(cudaEventRecord(dp_done_, dp_stream_));
    ACT_CUDA(cudaStreamWaitEvent(st, dp_done_, 0));
    if (dp_pack_running_) {
      dp_pack_running_ = false;
      ACT_CUDA(cudaStreamWaitEvent(st, dp_pack_done_, 0));
    }
    return 0;
  };
  if (!concurrent_) {
    if (backward(0, loss_grad, st)) return -1;
    if (p2p) {
      struct G0 { Engine* e; ~G0() { e->dp_fused_ = false; } } g0{this};
      dp_fused_ = true; dp_site_ = 0; dp_ranges_.clear();
      for (int r : {R_MAIN, R_CVAE, R_BACKBONE})
        if (region_n_[r] > 0 && (grads_ready(region_off_[r], region_n_[r], st) || adamw_region(r, step, st))) return -1;
      if (adamw_region(R_DEAD, step, st)) return -1;
      return p2p_finish();
    }
    if (cb) {                                               // serial data-parallel step: one averaging call per region
      for (int r : {R_MAIN, R_CVAE, R_BACKBONE})
        if (region_n_[r] > 0 && cb(cb_user, region_off_[r], region_n_[r], st)) { set_error("gradient all-reduce callback failed"); return -1; }
    }
    return adamw(step, st, false);
  }
  struct Guard {                                            // the hook is only live inside this call
    Engine* e;
    ~Guard() { e->grads_cb_ = nullptr; e->grads_cb_user_ = nullptr; e->inline_opt_on_ = false; e->inline_opt_step_ = 0; e->dp_bb_pending_ = false; e->dp_fused_ = false; e->dp_inline_from_ = 0; e->dp_defer_ = e->dp_defer_now_ = false; e->dp_deferred_.clear(); tc_set_sm_reserve(0); }
  } guard{this};
  grads_cb_ = cb; grads_cb_user_ = cb_user; opt_step_ = step;
  dp_fused_ = p2p; dp_site_ = 0; if (p2p) dp_ranges_.clear();
  dp_deferred_.clear();
  { static int defer = -1; if (defer < 0) { const char* e = getenv("ACT_DP_DEFER"); defer = (e && e[0] == '1') ? 1 : 0; }
    // opt-in (measured neutral: 3.408 vs 3.406 ms/step at N=2, 3.58 vs 3.56 at N=8): with one camera the backbone and the
    // CVAE branch end within ~50 us of each other, so their last ranges can share one barrier and run side by side.
    // With several cameras the backbone branch is the long one and the CVAE ranges are better exchanged under it.
    dp_defer_ = p2p && defer != 0 && cfg_.num_cameras == 1 && !cfg_.no_encoder; }
  static int inline_layers = -2;
  if (inline_layers == -2) { const char* e = getenv("ACT_DP_INLINE_LAYERS"); inline_layers = e ? atoi(e) : -1; }
  // peer-memory exchange: only the first `enc_layers / 2` CVAE layers of the backward (the highest-numbered ones) are
  // exchanged one by one; the rest form one range at the end of the branch
  dp_inline_from_ = p2p ? std::max(0, cfg_.enc_layers - (inline_layers >= 0 ? inline_layers : cfg_.enc_layers / 2)) : 0;
  const bool dp = cb != nullptr || p2p;                     // either way: ranges are exchanged in the order enqueued here
  if (wg_join(st)) return -1;                               // an asynchronous zero_grad
  if (zero_backward_outputs(st)) return -1;
  if (transformer_bwd(loss_grad, st)) return -1;
  auto main_regions = [&]() -> int {
    // MAIN / DEAD regions: every gradient and every read of these weights is ordered before this point of lane 0
    cudaStream_t wg = wg_gather(st, NWG);
    if (grads_ready(region_off_[R_MAIN], region_n_[R_MAIN], wg)) return -1;
    return adamw_region(R_MAIN, step, wg, 3) || adamw_region(R_DEAD, step, wg, 3) ? -1 : 0;
  };
  // With the hook the collectives run in the order they are enqueued here, so the region that is final first goes first
  if (dp && main_regions()) return -1;
  if (fork_side(st)) return -1;
  auto cvae_branch = [&]() -> int {
    if (cfg_.no_encoder) return 0;                          // encoder=None: no CVAE branch, region R_CVAE is empty
    inline_opt_step_ = step; inline_opt_on_ = true;     // (step may be 0 = "the device counter decides": a flag, not the step, gates the updates)
    const int rc = cvae_bwd(side_);
    inline_opt_step_ = 0; inline_opt_on_ = false;
    if (rc) return -1;
    // CVAE region: the encoder layers were updated inside cvae_bwd, the token projections / latent_proj follow here
    cudaStream_t wg = wg_gather(side_, NWG);
    const int nl = std::min(cfg_.enc_layers, 16);
    const long long tail = cvae_layer_off_[nl - 1] + cvae_layer_n_[nl - 1];
    // layers below dp_inline_from_ were not updated inside cvae_bwd (peer-memory exchange: the last layers' ranges would
    // only queue up behind each other on the serial exchange stream): they go with the head as ONE range
    const int lf = std::min(dp_inline_from_, nl);
    const long long head_end = lf > 0 ? cvae_layer_off_[lf - 1] + cvae_layer_n_[lf - 1] : cvae_layer_off_[0];
    const long long head_n = head_end - region_off_[R_CVAE];
    const long long tail_n = region_off_[R_CVAE] + region_n_[R_CVAE] - tail;
    const int bps = p2p ? 8 : 6;                            // nothing left to run beside these: whole chip
    if (dp_defer_) {                                        // exchanged with the backbone behind one barrier (p2p_finish)
      dp_defer_now_ = true;
      const int rc2 = adamw_range(region_off_[R_CVAE], head_n, cfg_.lr, true, step, wg, bps) ||
                      adamw_range(tail, tail_n, cfg_.lr, true, step, wg, bps);
      dp_defer_now_ = false;
      return rc2 ? -1 : 0;
    }
    if (grads_ready(region_off_[R_CVAE], head_n, wg)) return -1;
    if (adamw_range(region_off_[R_CVAE], head_n, cfg_.lr, true, step, wg, bps)) return -1;
    if ((!p2p || head_n <= 0) && grads_ready(tail, tail_n, wg)) return -1;   // (peer memory: the head's barrier covers the tail)
    return adamw_range(tail, tail_n, cfg_.lr, true, step, wg, bps);
  };
  static int bb_slot = -2;
  if (bb_slot == -2) { const char* e = getenv("ACT_DP_BB_SLOT"); bb_slot = e ? atoi(e) : -1; }
  dp_main_st_ = st;
  if (dp) {
    // backbone gradients are final roughly when half of the CVAE-encoder layers are (measured: profiles/r01_dp_timeline.md)
    dp_bb_slot_ = bb_slot >= 0 ? bb_slot : cfg_.enc_layers / 2;
    dp_bb_pending_ = true;
    if (cvae_branch()) return -1;
    if (dp_bb_pending_ && backbone_bwd_and_step(st)) return -1;
  } else {
    // the MAIN update is enqueued after the CVAE branch here: its (long) AdamW then does not sit ahead of the CVAE
    // branch's first kernels in the launch order
    if (cvae_branch()) return -1;
    if (main_regions()) return -1;
    if (backbone_bwd_and_step(st)) return -1;
  }
  if (join_side(st)) return -1;
  if (wg_join(st)) return -1;
  return p2p ? p2p_finish() : 0;
}

int Engine::backward(int stage, float loss_grad, cudaStream_t st) {
  ACT_CHECK(cfg_.train && grads_ != nullptr, "engine has no training workspace / gradient buffer");
  ACT_CHECK(have_actions_ && cur_B_ > 0, "backward needs a preceding training forward (with actions)");
  if (wg_join(st)) return -1;                               // an asynchronous zero_grad
  if (zero_backward_outputs(st)) return -1;
  if (stage == 0 || stage == 1) { if (transformer_bwd(loss_grad, st)) return -1; }
  // After the transformer the CVAE-encoder branch (needs dz) and the backbone branch (needs d_feat) share no buffers:
  // stage 0 / 4 run them as two parallel branches (side stream), stages 2 and 3 one at a time on the caller's stream.
  const bool both = stage == 0 || stage == 4;
  cudaStream_t sc = st;
  if (both) {
    if (fork_side(st)) return -1;
    if (concurrent_) sc = side_;
  }
  if ((both || stage == 2) && !cfg_.no_encoder) { if (cvae_bwd(sc)) return -1; }
  if (both || stage == 3) {
    cgy_off_ = 0;
    for (int k = 0; k < cfg_.num_cameras; ++k)
      if (backbone_bwd(k, cur_B_, st)) return -1;
  }
  if (sc != st && join_side(st)) return -1;
  if (wg_join(st)) return -1;
  return 0;
}

}  // namespace act
