// extern "C" surface of libact_b200.so (declared in include/act_b200.h).
#include "../../include/act_b200.h"
#include "tc_api.h"
#include "conv_ops.h"
#include "kernels.h"
#include "engine.h"
#include <cstring>

using namespace act;

extern "C" const char* act_last_error(void) { return last_error(); }
extern "C" int act_abi_version(void) { return 1; }

static EpiDesc to_epi(const act_epilogue* e) {
  EpiDesc d;
  if (e == nullptr) { d.flags = 0; return d; }      // rejected by check_epi ("epilogue without output")
  d.flags = e->flags; d.alpha = e->alpha; d.bias = e->bias; d.scale = e->scale; d.res_f32 = e->res_f32;
  d.res_bf16 = static_cast<const bf16*>(e->res_bf16); d.aux_bf16 = static_cast<const bf16*>(e->aux_bf16);
  d.out_f32 = e->out_f32; d.out_bf16 = static_cast<bf16*>(e->out_bf16);
  d.ldc = e->ldc; d.ldr = e->ldr; d.osz0 = e->osz0; d.osz1 = e->osz1; d.rsz0 = e->rsz0; d.rsz1 = e->rsz1;
  return d;
}

extern "C" int act_tc_gemm(int M, int N, int K, int nz0, int nz1, const void* A, int a_mn, int64_t lda, int64_t a_sz0,
                           int64_t a_sz1, const void* B, int b_mn, int64_t ldb, int64_t b_sz0, int64_t b_sz1,
                           const act_epilogue* epi, int ksplit, int bn_hint, int stages_hint, void* stream) {
  MatDesc a, b;
  a.ptr = static_cast<const bf16*>(A); a.mn_major = a_mn; a.ld = lda; a.sz0 = a_sz0; a.sz1 = a_sz1;
  b.ptr = static_cast<const bf16*>(B); b.mn_major = b_mn; b.ld = ldb; b.sz0 = b_sz0; b.sz1 = b_sz1;
  return tc_gemm(M, N, K, nz0, nz1, a, b, to_epi(epi), ksplit, static_cast<cudaStream_t>(stream), bn_hint, stages_hint);
}

#define ST(s) static_cast<cudaStream_t>(s)
#define BF(p) static_cast<const bf16*>(p)

extern "C" int act_cast_bf16(const float* src, void* dst, int64_t n, void* stream) {
  return cast_bf16(src, static_cast<bf16*>(dst), n, ST(stream));
}
extern "C" int act_pack_conv_weights(const float* w, const float* scale, void* wpk, void* wpkt, int Cout, int Cin, int KK,
                                     void* stream) {
  return pack_conv_weights(w, scale, static_cast<bf16*>(wpk), static_cast<bf16*>(wpkt), Cout, Cin, KK, ST(stream));
}
extern "C" int act_conv_fwd(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x,
                            const void* wpk, const act_epilogue* epi, void* stream) {
  ConvShape s{N, H, W, Cin, Cout, KH, stride, pad};
  return conv_fwd(s, BF(x), BF(wpk), to_epi(epi), ST(stream));
}
extern "C" int act_conv_dgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* dy,
                              const void* wpkt, const act_epilogue* epi, const void* res_compact, void* stream) {
  ConvShape s{N, H, W, Cin, Cout, KH, stride, pad};
  return conv_dgrad(s, BF(dy), BF(wpkt), to_epi(epi), BF(res_compact), ST(stream));
}
extern "C" int act_conv_wgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x,
                              const void* dy, float* dw, float alpha, const float* scale, void* stream) {
  ConvShape s{N, H, W, Cin, Cout, KH, stride, pad};
  return conv_wgrad(s, BF(x), BF(dy), dw, alpha, scale, ST(stream));
}
extern "C" int act_conv_wgrad_noalloc(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x,
                                      const void* dy, float* dw, float alpha, const float* scale, void* stream) {
  ConvShape s{N, H, W, Cin, Cout, KH, stride, pad};
  return conv_wgrad_noalloc(s, BF(x), BF(dy), dw, alpha, scale, ST(stream));
}

#define BFM(p) static_cast<bf16*>(p)
extern "C" int act_ln_fwd(const float* t, const float* gamma, const float* beta, const float* pos, int L, float* y,
                          void* yb, void* ypb, float* mean, float* rstd, int M, int D, void* stream) {
  return ln_fwd(t, gamma, beta, pos, L, y, BFM(yb), BFM(ypb), mean, rstd, M, D, ST(stream));
}
extern "C" int act_ln_bwd(const float* dy, const float* dy2, const float* t, const float* mean, const float* rstd,
                          const float* gamma, float* dt, void* dtb, float* dgamma, float* dbeta, int M, int D, void* stream) {
  return ln_bwd(dy, dy2, t, mean, rstd, gamma, dt, BFM(dtb), dgamma, dbeta, M, D, ST(stream));
}
extern "C" int act_add_pos_cast(const float* x, const float* pos, int L, void* xb, void* xpb, int M, int D, void* stream) {
  return add_pos_cast(x, pos, L, BFM(xb), BFM(xpb), M, D, ST(stream));
}
extern "C" int act_softmax_fwd(const float* S, const uint8_t* mask, void* P, int R, int L, int Lp, int rpb, void* stream) {
  return softmax_fwd(S, mask, BFM(P), R, L, Lp, rpb, ST(stream));
}
extern "C" int act_softmax_bwd(const void* P, const float* dP, void* dS, int R, int L, int Lp, float scale, void* stream) {
  return softmax_bwd(BF(P), dP, BFM(dS), R, L, Lp, scale, ST(stream));
}
extern "C" int act_attn_scores(int bwd, const void* A, int64_t lda, int64_t a_sz0, int64_t a_sz1, const void* Bm, int64_t ldb,
                               int64_t b_sz0, int64_t b_sz1, int Lq, int Lk, int Lkp, int heads, int batch, float scale,
                               const uint8_t* mask, const void* P, void* out, void* stream) {
  MatDesc a, b;
  a.ptr = BF(A); a.ld = lda; a.sz0 = a_sz0; a.sz1 = a_sz1;
  b.ptr = BF(Bm); b.ld = ldb; b.sz0 = b_sz0; b.sz1 = b_sz1;
  return tc_attn_scores(bwd, a, b, Lq, Lk, Lkp, heads, batch, scale, mask, BF(P), BFM(out), ST(stream));
}
extern "C" int act_attn_scores_bwd(const void* dO, int64_t ld_do, int64_t do_sz0, int64_t do_sz1, const void* V, int64_t ldv,
                                   int64_t v_sz0, int64_t v_sz1, const void* O, int64_t ldo, int64_t o_sz0, int64_t o_sz1, int Lq,
                                   int Lk, int Lkp, int heads, int batch, float scale, const void* P, void* dS, void* stream) {
  MatDesc a, b, o;
  a.ptr = BF(dO); a.ld = ld_do; a.sz0 = do_sz0; a.sz1 = do_sz1;
  b.ptr = BF(V); b.ld = ldv; b.sz0 = v_sz0; b.sz1 = v_sz1;
  o.ptr = BF(O); o.ld = ldo; o.sz0 = o_sz0; o.sz1 = o_sz1;
  return tc_attn_scores(1, a, b, Lq, Lk, Lkp, heads, batch, scale, nullptr, BF(P), BFM(dS), ST(stream), nullptr, nullptr, O ? &o : nullptr);
}
extern "C" int act_colsum_bf16(const void* X, int64_t ld, int M, int N, float* out, void* stream) {
  return colsum_bf16(BF(X), ld, M, N, out, ST(stream));
}
static IdxMap to_map(const int64_t* m) { IdxMap r; r.rpb = static_cast<int>(m[0]); r.bs = m[1]; r.s = m[2]; return r; }
extern "C" int act_sgemm(int I, int J, int R, const float* A, const int64_t* a_i, const int64_t* a_r, const float* B,
                         const int64_t* b_r, const int64_t* b_j, float* C, const int64_t* c_i, const int64_t* c_j,
                         const float* bias, float alpha, int accumulate, void* stream) {
  SgemmArgs a;
  a.I = I; a.J = J; a.R = R; a.A = A; a.ai = to_map(a_i); a.ar = to_map(a_r); a.B = B; a.br = to_map(b_r);
  a.bj = to_map(b_j); a.C = C; a.ci = to_map(c_i); a.cj = to_map(c_j); a.bias = bias; a.alpha = alpha; a.accumulate = accumulate;
  return sgemm_generic(a, ST(stream));
}
extern "C" int act_loss(const float* a_hat, const float* actions, int64_t abs_, const uint8_t* is_pad, int64_t pbs,
                        const float* mu, const float* logvar, int B, int nq, int A, int Z, float klw, float gscale,
                        float* losses, float* d_a_hat, float* d_mu, float* d_logvar, void* stream) {
  return act::act_loss(a_hat, actions, abs_, is_pad, pbs, mu, logvar, B, nq, A, Z, klw, gscale, losses, d_a_hat, d_mu,
                       d_logvar, ST(stream));
}
extern "C" int act_reparam_fwd(const float* info, const float* eps, float* mu, float* logvar, float* z, int B, int Z, void* stream) {
  return reparam_fwd(info, eps, mu, logvar, z, B, Z, ST(stream));
}
extern "C" int act_reparam_bwd(const float* dz, const float* eps, const float* logvar, const float* dmu, const float* dlv,
                               float* d_info, int B, int Z, void* stream) {
  return reparam_bwd(dz, eps, logvar, dmu, dlv, d_info, B, Z, ST(stream));
}
extern "C" int act_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, void* stream) {
  return maxpool_fwd(BF(x), BFM(y), idx, N, H, W, C, ST(stream));
}
extern "C" int act_maxpool_bwd(const void* dy, const uint8_t* idx, const void* x, void* dx, int N, int H, int W, int C, void* stream) {
  return maxpool_bwd(BF(dy), idx, BF(x), BFM(dx), N, H, W, C, ST(stream));
}
extern "C" int act_adamw_step(float* p, const float* g, float* m, float* v, void* shadow, int64_t n, float lr, float b1,
                              float b2, float eps, float wd, int step, int mode, void* stream) {
  AdamWArgs a;
  a.p = p; a.g = g; a.m = m; a.v = v; a.shadow = BFM(shadow); a.n = n; a.lr = lr; a.beta1 = b1; a.beta2 = b2; a.eps = eps;
  a.wd = wd; a.step = step; a.mode = mode;
  return adamw_step(a, ST(stream));
}
extern "C" int act_temporal_agg_step(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A,
                                     double k, void* stream) {
  return temporal_agg_step(ring, chunks, t, out, E, nq, A, k, ST(stream));
}

extern "C" int act_temporal_agg_step_ex(float* ring, const float* chunks, const int* t, double* out, int E, int nq, int A,
                                        double k, const float* action_mean, const float* action_std, void* stream) {
  return temporal_agg_step(ring, chunks, t, out, E, nq, A, k, ST(stream), action_mean, action_std);
}
extern "C" int act_chunk_select(const float* chunks, const int* t, double* out, int E, int nq, int A, int base_delay,
                                const float* action_mean, const float* action_std, void* stream) {
  return chunk_select(chunks, t, out, E, nq, A, base_delay, action_mean, action_std, ST(stream));
}

struct act_engine { Engine* eng; };

extern "C" int act_engine_create(const act_engine_config* c, act_engine** out) {
  if (!c || !out) { set_error("null argument"); return -1; }
  EngineConfig k;
  k.batch_max = c->batch_max; k.num_cameras = c->num_cameras; k.img_h = c->img_h; k.img_w = c->img_w;
  k.hidden_dim = c->hidden_dim; k.nheads = c->nheads; k.dim_ff = c->dim_feedforward; k.enc_layers = c->enc_layers;
  k.dec_layers = c->dec_layers; k.num_queries = c->num_queries; k.state_dim = c->state_dim; k.action_dim = c->action_dim;
  k.latent_dim = c->latent_dim; k.kl_weight = c->kl_weight; k.lr = c->lr; k.lr_backbone = c->lr_backbone;
  k.weight_decay = c->weight_decay; k.beta1 = c->beta1; k.beta2 = c->beta2; k.adam_eps = c->adam_eps;
  k.no_encoder = c->no_encoder; k.train = c->train;
  Engine* e = new Engine(k);
  if (e->init() != 0) { delete e; return -1; }
  *out = new act_engine{e};
  return 0;
}
extern "C" void act_engine_destroy(act_engine* e) { if (e) { delete e->eng; delete e; } }
extern "C" int act_engine_num_tensors(const act_engine* e) { return static_cast<int>(e->eng->tensors().size()); }
extern "C" int64_t act_engine_total_floats(const act_engine* e) { return e->eng->total_floats(); }
extern "C" int64_t act_engine_workspace_bytes(const act_engine* e) { return e->eng->workspace_bytes(); }
extern "C" int act_engine_tensor_info(const act_engine* e, int i, char* name, int name_cap, int64_t shape[4], int* ndim,
                                      int64_t* offset, int64_t* numel, int* kind, int* region) {
  const auto& ts = e->eng->tensors();
  if (i < 0 || i >= static_cast<int>(ts.size())) { set_error("tensor index out of range"); return -1; }
  const TensorEntry& t = ts[i];
  if (name && name_cap > 0) { strncpy(name, t.name.c_str(), name_cap - 1); name[name_cap - 1] = 0; }
  for (int k = 0; k < 4; ++k) shape[k] = t.shape[k];
  *ndim = t.ndim; *offset = t.offset; *numel = t.numel; *kind = t.kind; *region = t.region;
  return 0;
}
extern "C" int act_engine_region(const act_engine* e, int region, int64_t* offset, int64_t* count) {
  if (region < 0 || region >= R_COUNT) { set_error("bad region"); return -1; }
  long long o, n;
  e->eng->region(region, &o, &n);
  *offset = o; *count = n;
  return 0;
}
extern "C" int act_engine_bind(act_engine* e, float* params, float* grads, float* m, float* v) { return e->eng->bind(params, grads, m, v); }
extern "C" int act_engine_refresh(act_engine* e, void* stream) { return e->eng->refresh_shadows(ST(stream)); }
extern "C" int act_engine_forward(act_engine* e, int B, const float* qpos, const float* image, const float* actions,
                                  int64_t actions_bs, const uint8_t* is_pad, int64_t is_pad_bs, const float* eps,
                                  float* a_hat, float* is_pad_hat, float* mu, float* logvar, float* losses, void* stream) {
  return e->eng->forward(B, qpos, image, actions, actions_bs, is_pad, is_pad_bs, eps, a_hat, is_pad_hat, mu, logvar, losses, ST(stream));
}
extern "C" int act_engine_backward(act_engine* e, int stage, float loss_grad, void* stream) { return e->eng->backward(stage, loss_grad, ST(stream)); }
extern "C" int act_engine_zero_grad(act_engine* e, void* stream) { return e->eng->zero_grad(ST(stream)); }
extern "C" int act_engine_zero_grad_async(act_engine* e, void* stream) { return e->eng->zero_grad(ST(stream), true); }
extern "C" int act_engine_backward_and_step(act_engine* e, int step, float loss_grad, void* stream) {
  return e->eng->backward_and_step(step, loss_grad, ST(stream));
}
extern "C" int act_engine_backward_and_step_dp(act_engine* e, int step, float loss_grad, void* stream,
                                               act_grads_ready_fn grads_ready, void* user) {
  if (!grads_ready) { set_error("act_engine_backward_and_step_dp: null callback"); return -1; }
  return e->eng->backward_and_step(step, loss_grad, ST(stream), grads_ready, user);
}
extern "C" int act_engine_adamw_step(act_engine* e, int step, void* stream) { return e->eng->adamw(step, ST(stream)); }
extern "C" int act_engine_dp_local(act_engine* e, void** grads, void** params, void** shadow, void** signal) {
  return e->eng->dp_local(grads, params, shadow, signal);
}
extern "C" int act_engine_dp_export(act_engine* e, void* blob, int capacity) { return e->eng->dp_export(blob, capacity); }
extern "C" int act_engine_dp_connect(act_engine* e, int rank, int world, const void* blobs, int blob_bytes) {
  return e->eng->dp_connect(rank, world, blobs, blob_bytes);
}
extern "C" int act_engine_dp_set_peers(act_engine* e, int rank, int world, void* const* grads, void* const* params,
                                       void* const* shadow, void* const* signal) {
  return e->eng->dp_set_peers(rank, world, grads, params, shadow, signal);
}
extern "C" int act_engine_dp_use_buffers(act_engine* e, void* shadow, void* signal) { return e->eng->dp_use_buffers(shadow, signal); }
extern "C" int64_t act_engine_dp_signal_bytes(void) { return Engine::DP_SIGNAL_BYTES; }
extern "C" int act_engine_dp_set_multicast(act_engine* e, void* mc_grads, void* mc_params, void* mc_shadow) {
  return e->eng->dp_set_multicast(mc_grads, mc_params, mc_shadow);
}
extern "C" int act_multimem_reduce_f32(const float* mc_src, float* dst_local, int64_t n, void* stream) {
  return multimem_reduce_f32(mc_src, dst_local, n, ST(stream));
}
extern "C" int act_multimem_bcast_f32(float* mc_dst, const float* src_local, int64_t n, void* stream) {
  return multimem_bcast_f32(mc_dst, src_local, n, ST(stream));
}
extern "C" int act_engine_backward_and_step_p2p(act_engine* e, int step, float loss_grad, void* stream) {
  return e->eng->backward_and_step(step, loss_grad, ST(stream), nullptr, nullptr, true);
}
extern "C" int act_engine_dp_status(act_engine* e) { return e->eng->dp_status(); }
extern "C" int act_engine_dp_preset_signal(act_engine* e, int peer_rank, unsigned int epoch) {
  return e->eng->dp_preset_signal(peer_rank, epoch);
}
extern "C" int act_engine_dp_slices(act_engine* e, long long* offsets, long long* counts, int capacity) {
  return e->eng->dp_slices(offsets, counts, capacity);
}
extern "C" int act_engine_adamw_region(act_engine* e, int step, int region, int blocks_per_sm, void* stream) {
  if (region < 0 || region >= R_COUNT) { set_error("act_engine_adamw_region: bad region"); return -1; }
  if (step <= 0) { set_error("act_engine_adamw_region: step must be >= 1"); return -1; }
  if (e->eng->begin_opt_step(step, ST(stream))) return -1;      // idempotent: every region of one step sets the same value
  return e->eng->adamw_region(region, step, ST(stream), blocks_per_sm);
}
extern "C" int act_engine_set_image_format(act_engine* e, int fmt) { return e->eng->set_image_format(fmt); }
extern "C" int act_engine_set_concurrent(act_engine* e, int on) { return e->eng->set_concurrent(on != 0); }
extern "C" int act_engine_set_lr(act_engine* e, float lr, float lr_backbone) { e->eng->set_lr(lr, lr_backbone); return 0; }
extern "C" int act_engine_set_qpos_norm(act_engine* e, const float* mean, const float* std, void* stream) {
  return e->eng->set_qpos_norm(mean, std, ST(stream));
}
extern "C" int act_engine_enable_qpos_norm(act_engine* e, int on) {
  if (on && !e->eng->qnorm_set_) { set_error("act_engine_enable_qpos_norm: no statistics stored"); return -1; }
  e->eng->qnorm_on_ = on != 0;
  return 0;
}
extern "C" int act_engine_set_dropout(act_engine* e, float p, uint64_t seed) { return e->eng->set_dropout(p, seed); }
extern "C" int act_engine_dropout_state(act_engine* e, float* p, uint64_t* seed, uint32_t* step) {
  unsigned long long s = 0;
  const int rc = e->eng->dropout_state(p, &s, step);
  if (seed) *seed = s;
  return rc;
}
extern "C" int act_dropout_mask(uint64_t seed, uint32_t site, uint32_t step, float p, int64_t n, uint8_t* out, void* stream) {
  return dropout_mask(seed, site, step, p, n, out, ST(stream));
}
extern "C" int act_engine_set_loss_grad(act_engine* e, const float* gvec3) { e->eng->set_loss_grad(gvec3); return 0; }
extern "C" int act_engine_join_lanes(act_engine* e, void* stream) { return e->eng->join_lanes(ST(stream)); }
extern "C" int act_engine_launches(act_engine* e, int reset) { const int n = e->eng->launches(); if (reset) e->eng->reset_launches(); return n; }

extern "C" int act_engine_debug_buffer(act_engine* e, const char* name, void* dst, int64_t capacity, int64_t* numel,
                                       int* dtype, void* stream) {
  long long n = 0;
  void* src = nullptr;
  int dt = 0;
  if (e->eng->debug_buffer(name, &src, &n, &dt)) return -1;
  if (numel) *numel = n;
  if (dtype) *dtype = dt;
  if (!dst) return 0;
  const long long bytes = n * (dt == 0 ? 4 : 2);
  if (capacity < bytes) { set_error("act_engine_debug_buffer: destination too small"); return -1; }
  if (cudaMemcpyAsync(dst, src, static_cast<size_t>(bytes), cudaMemcpyDeviceToDevice, ST(stream)) != cudaSuccess) {
    set_error("act_engine_debug_buffer: copy failed"); return -1;
  }
  return 0;
}

extern "C" int act_tc_set_skip_mask(uint32_t mask) { tc_set_skip_mask(mask); return 0; }
extern "C" double act_tc_skipped_flops(void) { return tc_skipped_flops(); }
extern "C" int act_tc_debug_timestamps(void* buf) { tc_debug_timestamps(static_cast<unsigned long long*>(buf)); return 0; }
extern "C" int act_tc_profile_begin(void) { tc_profile_begin(); return 0; }
extern "C" int act_tc_profile_end(double* total_ms, double* total_flops, int* launches, double* max_ms) {
  return tc_profile_end(total_ms, total_flops, launches, max_ms);
}
extern "C" int act_tc_profile_set_dump(const char* path) { tc_profile_set_dump(path); return 0; }
