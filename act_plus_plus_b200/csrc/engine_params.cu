// Engine: tensor table (reference state_dict names), flat layout by optimizer region, bf16 shadows, AdamW.
#include "engine.h"
#include <cuda.h>
#include <cmath>
#include <cstring>

namespace act {

Engine::Engine(const EngineConfig& cfg) : cfg_(cfg) {}

Engine::~Engine() {
  for (void* p : dp_opened_) cudaIpcCloseMemHandle(p);
  for (cudaEvent_t e : dp_ev_) if (e) cudaEventDestroy(e);
  if (dp_done_) cudaEventDestroy(dp_done_);
  if (dp_pack_ev_) cudaEventDestroy(dp_pack_ev_);
  if (dp_pack_done_) cudaEventDestroy(dp_pack_done_);
  if (dp_pack_stream_) cudaStreamDestroy(dp_pack_stream_);
  if (dp_stream_) cudaStreamDestroy(dp_stream_);
  for (void* p : allocs_) cudaFree(p);
  if (ev_fork_) cudaEventDestroy(ev_fork_);
  if (ev_join_) cudaEventDestroy(ev_join_);
  if (side_) cudaStreamDestroy(side_);
  if (ev_stage_) cudaEventDestroy(ev_stage_);
  if (ev_zero_f_) cudaEventDestroy(ev_zero_f_);
  if (ev_zero_b_) cudaEventDestroy(ev_zero_b_);
  for (int b = 0; b < 2; ++b)
    for (int l = 0; l < 2; ++l) {
      if (aux_ev_[b][l]) cudaEventDestroy(aux_ev_[b][l]);
      if (aux_done_[b][l]) cudaEventDestroy(aux_done_[b][l]);
      if (aux_[b][l]) cudaStreamDestroy(aux_[b][l]);
    }
  for (int b = 0; b < 2; ++b) {
    if (wg_ev_[b]) cudaEventDestroy(wg_ev_[b]);
    for (int i = 0; i <= NWG; ++i) {
      if (wg_done_[b][i]) cudaEventDestroy(wg_done_[b][i]);
      if (wg_[b][i]) cudaStreamDestroy(wg_[b][i]);
    }
  }
}

int Engine::ensure_streams() {
  if (side_) return 0;
  const char* env = getenv("ACT_CONCURRENT");
  concurrent_ = !(env && env[0] == '0');
  if (const char* r = getenv("ACT_SM_RESERVE")) sm_reserve_ = atoi(r);
  int prio_lo = 0, prio_hi = 0;
  ACT_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  const char* pe = getenv("ACT_SIDE_PRIO");
  // ACT_SIDE_PRIO=1 gives the CVAE branch (a long chain of small kernels) SM slots ahead of the backbone branch; measured
  // neutral-to-worse (the persistent conv kernels then wait for their late CTAs), so the default is equal priority
  ACT_CUDA(cudaStreamCreateWithPriority(&side_, cudaStreamNonBlocking, (pe && pe[0] == '1') ? prio_hi : prio_lo));
  ACT_CUDA(cudaEventCreateWithFlags(&ev_fork_, cudaEventDisableTiming));
  ACT_CUDA(cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming));
  ACT_CUDA(cudaEventCreateWithFlags(&ev_stage_, cudaEventDisableTiming));
  ACT_CUDA(cudaEventCreateWithFlags(&ev_zero_f_, cudaEventDisableTiming));
  ACT_CUDA(cudaEventCreateWithFlags(&ev_zero_b_, cudaEventDisableTiming));
  for (int b = 0; b < 2; ++b)
    for (int l = 0; l < 2; ++l) {
      ACT_CUDA(cudaStreamCreateWithFlags(&aux_[b][l], cudaStreamNonBlocking));
      ACT_CUDA(cudaEventCreateWithFlags(&aux_ev_[b][l], cudaEventDisableTiming));
      ACT_CUDA(cudaEventCreateWithFlags(&aux_done_[b][l], cudaEventDisableTiming));
    }
  for (int b = 0; b < 2; ++b) {
    ACT_CUDA(cudaEventCreateWithFlags(&wg_ev_[b], cudaEventDisableTiming));
    for (int i = 0; i <= NWG; ++i) {
      ACT_CUDA(cudaStreamCreateWithFlags(&wg_[b][i], cudaStreamNonBlocking));
      ACT_CUDA(cudaEventCreateWithFlags(&wg_done_[b][i], cudaEventDisableTiming));
    }
  }
  return 0;
}
int Engine::fork_side(cudaStream_t st) {
  if (ensure_streams()) return -1;
  ACT_CUDA(cudaEventRecord(ev_fork_, st));
  ACT_CUDA(cudaStreamWaitEvent(side_, ev_fork_, 0));
  return 0;
}
cudaStream_t Engine::aux_fork(cudaStream_t st, int lane) {
  static int aux_on = -1;
  if (aux_on < 0) { const char* e = getenv("ACT_AUX"); aux_on = !(e && e[0] == '0'); }
  if (ensure_streams() || !concurrent_ || !aux_on) return st;
  const int b = (st == side_) ? 1 : 0;
  if (cudaEventRecord(aux_ev_[b][lane], st) != cudaSuccess || cudaStreamWaitEvent(aux_[b][lane], aux_ev_[b][lane], 0) != cudaSuccess) {
    set_error("aux lane fork failed");
    return st;
  }
  aux_used_[b][lane] = true;
  return aux_[b][lane];
}
int Engine::aux_join(cudaStream_t st, int lane) {
  if (!side_ || !concurrent_) return 0;
  const int b = (st == side_) ? 1 : 0;
  if (!aux_used_[b][lane]) return 0;
  aux_used_[b][lane] = false;
  ACT_CUDA(cudaEventRecord(aux_done_[b][lane], aux_[b][lane]));
  ACT_CUDA(cudaStreamWaitEvent(st, aux_done_[b][lane], 0));
  return 0;
}
cudaStream_t Engine::wg_fork(cudaStream_t st, int lane) {
  if (ensure_streams() || !concurrent_) return st;
  const int b = (st == side_) ? 1 : 0;
  const int i = lane >= 0 ? lane : (wg_rr_[b]++ % NWG);
  if (cudaEventRecord(wg_ev_[b], st) != cudaSuccess || cudaStreamWaitEvent(wg_[b][i], wg_ev_[b], 0) != cudaSuccess) {
    set_error("weight-gradient stream fork failed");
    return st;
  }
  wg_used_[b][i] = true;
  return wg_[b][i];
}
cudaStream_t Engine::wg_gather(cudaStream_t st, int lane) {
  if (ensure_streams() || !concurrent_) return st;
  const int b = (st == side_) ? 1 : 0;
  for (int i = 0; i <= NWG; ++i) {
    if (i == lane || !wg_used_[b][i]) continue;
    if (cudaEventRecord(wg_done_[b][i], wg_[b][i]) != cudaSuccess || cudaStreamWaitEvent(wg_[b][lane], wg_done_[b][i], 0) != cudaSuccess)
      set_error("weight-gradient stream gather failed");
  }
  return wg_fork(st, lane);
}
int Engine::wg_join(cudaStream_t st) {
  for (int b = 0; b < 2; ++b)
    for (int i = 0; i <= NWG; ++i) {
      if (!wg_used_[b][i]) continue;
      ACT_CUDA(cudaEventRecord(wg_done_[b][i], wg_[b][i]));
      ACT_CUDA(cudaStreamWaitEvent(st, wg_done_[b][i], 0));
      wg_used_[b][i] = false;
    }
  return 0;
}
float* Engine::gz_alloc(long long elems) {
  elems = (elems + 31) & ~31LL;
  if (gz_off_ + elems > gz_cap_) { set_error("backward split-K output arena overflow"); return nullptr; }
  float* p = gz_arena_ + gz_off_;
  gz_off_ += elems;
  return p;
}
int Engine::zero_backward_outputs(cudaStream_t st) {
  if (!splitk_on() || !gz_arena_) return 0;
  cudaStream_t z = wg_fork(st, 1);
  ACT_CUDA(cudaMemsetAsync(gz_arena_, 0, sizeof(float) * gz_cap_, z));
  zero_b_pending_ = z != st;
  if (zero_b_pending_) ACT_CUDA(cudaEventRecord(ev_zero_b_, z));
  return 0;
}
bf16* Engine::cgy_alloc(long long elems) {
  elems = (elems + 127) & ~127LL;
  if (cgy_off_ + elems > cgy_cap_) { set_error("backward conv-gradient arena overflow"); return nullptr; }
  bf16* p = cgy_arena_ + cgy_off_;
  cgy_off_ += elems;
  return p;
}
bf16* Engine::gy_alloc(long long elems) {
  elems = (elems + 127) & ~127LL;
  if (gy_off_ + elems > gy_lim_) { set_error("backward dY arena overflow"); return nullptr; }
  bf16* p = gy_arena_ + gy_off_;
  gy_off_ += elems;
  return p;
}
int Engine::join_side(cudaStream_t st) {
  ACT_CUDA(cudaEventRecord(ev_join_, side_));
  ACT_CUDA(cudaStreamWaitEvent(st, ev_join_, 0));
  return 0;
}

template <typename T>
T* Engine::alloc(long long n) {
  if (n <= 0) n = 1;
  void* p = nullptr;
  const size_t bytes = (static_cast<size_t>(n) * sizeof(T) + 255) & ~size_t(255);
  if (cudaMalloc(&p, bytes) != cudaSuccess) { set_error("cudaMalloc failed for " + std::to_string(bytes) + " bytes"); return nullptr; }
  cudaMemset(p, 0, bytes);
  allocs_.push_back(p);
  ws_bytes_ += bytes;
  return static_cast<T*>(p);
}
template float* Engine::alloc<float>(long long);
template bf16* Engine::alloc<bf16>(long long);
template uint8_t* Engine::alloc<uint8_t>(long long);

TensorEntry* Engine::find(const std::string& name) {
  auto it = index_.find(name);
  if (it == index_.end()) { set_error("unknown tensor " + name); return &tensors_[0]; }
  return &tensors_[it->second];
}

static const char* kStages[4] = {"layer1", "layer2", "layer3", "layer4"};
static const int kStageCin[4] = {64, 64, 128, 256}, kStageCout[4] = {64, 128, 256, 512}, kStageStride[4] = {1, 2, 2, 2};

// Same order as the reference's state_dict (detr_vae.py:38-105 registration order; oracle/act_oracle.py:state_spec).
void Engine::build_tensor_table() {
  const int d = cfg_.hidden_dim, ff = cfg_.dim_ff, nq = cfg_.num_queries, S = cfg_.state_dim, A = cfg_.action_dim, Z = cfg_.latent_dim;
  auto add = [&](const std::string& name, std::vector<long long> shape, int kind, int region) {
    TensorEntry t;
    t.name = name; t.ndim = static_cast<int>(shape.size()); t.numel = 1;
    for (size_t i = 0; i < shape.size(); ++i) { t.shape[i] = shape[i]; t.numel *= shape[i]; }
    t.kind = kind; t.region = region; t.order = static_cast<int>(tensors_.size());
    tensors_.push_back(t);
  };
  auto enc_layer = [&](const std::string& p, int kind, int region) {
    add(p + "self_attn.in_proj_weight", {3 * d, d}, kind, region); add(p + "self_attn.in_proj_bias", {3 * d}, kind, region);
    add(p + "self_attn.out_proj.weight", {d, d}, kind, region); add(p + "self_attn.out_proj.bias", {d}, kind, region);
    add(p + "linear1.weight", {ff, d}, kind, region); add(p + "linear1.bias", {ff}, kind, region);
    add(p + "linear2.weight", {d, ff}, kind, region); add(p + "linear2.bias", {d}, kind, region);
    add(p + "norm1.weight", {d}, kind, region); add(p + "norm1.bias", {d}, kind, region);
    add(p + "norm2.weight", {d}, kind, region); add(p + "norm2.bias", {d}, kind, region);
  };
  auto dec_layer = [&](const std::string& p, int kind, int region) {
    add(p + "self_attn.in_proj_weight", {3 * d, d}, kind, region); add(p + "self_attn.in_proj_bias", {3 * d}, kind, region);
    add(p + "self_attn.out_proj.weight", {d, d}, kind, region); add(p + "self_attn.out_proj.bias", {d}, kind, region);
    add(p + "multihead_attn.in_proj_weight", {3 * d, d}, kind, region); add(p + "multihead_attn.in_proj_bias", {3 * d}, kind, region);
    add(p + "multihead_attn.out_proj.weight", {d, d}, kind, region); add(p + "multihead_attn.out_proj.bias", {d}, kind, region);
    add(p + "linear1.weight", {ff, d}, kind, region); add(p + "linear1.bias", {ff}, kind, region);
    add(p + "linear2.weight", {d, ff}, kind, region); add(p + "linear2.bias", {d}, kind, region);
    for (int i = 1; i <= 3; ++i) {
      add(p + "norm" + std::to_string(i) + ".weight", {d}, kind, region);
      add(p + "norm" + std::to_string(i) + ".bias", {d}, kind, region);
    }
  };
  auto bn = [&](const std::string& p, int c) {
    add(p + "weight", {c}, T_BUFFER, R_BUFFER); add(p + "bias", {c}, T_BUFFER, R_BUFFER);
    add(p + "running_mean", {c}, T_BUFFER, R_BUFFER); add(p + "running_var", {c}, T_BUFFER, R_BUFFER);
  };
  add("pos_table", {1, nq + 2, d}, T_BUFFER, R_BUFFER);
  for (int i = 0; i < cfg_.enc_layers; ++i) enc_layer("transformer.encoder.layers." + std::to_string(i) + ".", T_PARAM, R_MAIN);
  for (int i = 0; i < cfg_.dec_layers; ++i)
    dec_layer("transformer.decoder.layers." + std::to_string(i) + ".", i == 0 ? T_PARAM : T_PARAM_DEAD, i == 0 ? R_MAIN : R_DEAD);
  add("transformer.decoder.norm.weight", {d}, T_PARAM, R_MAIN); add("transformer.decoder.norm.bias", {d}, T_PARAM, R_MAIN);
  // encoder=None (detr_vae.py:373-377): no CVAE-encoder stack; its token / latent projections still exist in the
  // state_dict but never take part in a forward, so autograd leaves their .grad None and AdamW skips them
  const bool cvae = cfg_.no_encoder == 0;
  const int ck = cvae ? T_PARAM : T_PARAM_NOGRAD, cr = cvae ? R_CVAE : R_NOGRAD;
  if (cvae) for (int i = 0; i < cfg_.enc_layers; ++i) enc_layer("encoder.layers." + std::to_string(i) + ".", T_PARAM, R_CVAE);
  add("action_head.weight", {A, d}, T_PARAM, R_MAIN); add("action_head.bias", {A}, T_PARAM, R_MAIN);
  add("is_pad_head.weight", {1, d}, T_PARAM_NOGRAD, R_NOGRAD); add("is_pad_head.bias", {1}, T_PARAM_NOGRAD, R_NOGRAD);
  add("query_embed.weight", {nq, d}, T_PARAM, R_MAIN);
  add("input_proj.weight", {d, 512, 1, 1}, T_PARAM, R_MAIN); add("input_proj.bias", {d}, T_PARAM, R_MAIN);
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    const std::string p = "backbones." + std::to_string(k) + ".0.body.";
    add(p + "conv1.weight", {64, 3, 7, 7}, T_PARAM_BACKBONE, R_BACKBONE);
    bn(p + "bn1.", 64);
    for (int s = 0; s < 4; ++s)
      for (int b = 0; b < 2; ++b) {
        const std::string q = p + kStages[s] + "." + std::to_string(b) + ".";
        const int cin = b == 0 ? kStageCin[s] : kStageCout[s], cout = kStageCout[s];
        add(q + "conv1.weight", {cout, cin, 3, 3}, T_PARAM_BACKBONE, R_BACKBONE); bn(q + "bn1.", cout);
        add(q + "conv2.weight", {cout, cout, 3, 3}, T_PARAM_BACKBONE, R_BACKBONE); bn(q + "bn2.", cout);
        if (b == 0 && (kStageStride[s] != 1 || cin != cout)) {
          add(q + "downsample.0.weight", {cout, cin, 1, 1}, T_PARAM_BACKBONE, R_BACKBONE); bn(q + "downsample.1.", cout);
        }
      }
  }
  add("input_proj_robot_state.weight", {d, S}, T_PARAM, R_MAIN); add("input_proj_robot_state.bias", {d}, T_PARAM, R_MAIN);
  add("cls_embed.weight", {1, d}, ck, cr);
  add("encoder_action_proj.weight", {d, A}, ck, cr); add("encoder_action_proj.bias", {d}, ck, cr);
  add("encoder_joint_proj.weight", {d, S}, ck, cr); add("encoder_joint_proj.bias", {d}, ck, cr);
  add("latent_proj.weight", {2 * Z, d}, ck, cr); add("latent_proj.bias", {2 * Z}, ck, cr);
  add("latent_out_proj.weight", {d, Z}, T_PARAM, R_MAIN); add("latent_out_proj.bias", {d}, T_PARAM, R_MAIN);
  add("additional_pos_embed.weight", {2, d}, T_PARAM, R_MAIN);

  // flat layout: regions in all-reduce order, tensors 8-float aligned (bf16 shadow stays 16-byte aligned for TMA)
  long long off = 0;
  const int order[R_COUNT] = {R_MAIN, R_CVAE, R_BACKBONE, R_DEAD, R_NOGRAD, R_BUFFER};
  for (int ri = 0; ri < R_COUNT; ++ri) {
    const int r = order[ri];
    region_off_[r] = off;
    for (auto& t : tensors_)
      if (t.region == r) { t.offset = off; off += (t.numel + 7) / 8 * 8; }
    region_n_[r] = off - region_off_[r];
  }
  total_floats_ = off;
  for (int l = 0; cvae && l < cfg_.enc_layers && l < 16; ++l) {     // contiguous span of CVAE-encoder layer l inside R_CVAE
    const std::string pre = "encoder.layers." + std::to_string(l) + ".";
    long long lo = -1, hi = -1;
    for (auto& t : tensors_)
      if (t.region == R_CVAE && t.name.compare(0, pre.size(), pre) == 0) {
        if (lo < 0 || t.offset < lo) lo = t.offset;
        hi = std::max(hi, t.offset + (t.numel + 7) / 8 * 8);
      }
    cvae_layer_off_[l] = lo; cvae_layer_n_[l] = hi - lo;
  }
  for (int k = 0; k < cfg_.num_cameras && k < 8; ++k) {     // camera k's span of R_BACKBONE, split at the first layer4 tensor
    const std::string pre = "backbones." + std::to_string(k) + ".0.body.", l4 = pre + "layer4.";
    long long lo = -1, hi = -1, l4lo = -1;
    for (auto& t : tensors_)
      if (t.region == R_BACKBONE && t.name.compare(0, pre.size(), pre) == 0) {
        if (lo < 0 || t.offset < lo) lo = t.offset;
        hi = std::max(hi, t.offset + (t.numel + 7) / 8 * 8);
        if (t.name.compare(0, l4.size(), l4) == 0 && (l4lo < 0 || t.offset < l4lo)) l4lo = t.offset;
      }
    bb_rest_off_[k] = lo; bb_rest_n_[k] = l4lo - lo; bb_l4_off_[k] = l4lo; bb_l4_n_[k] = hi - l4lo;
  }
  for (size_t i = 0; i < tensors_.size(); ++i) index_[tensors_[i].name] = static_cast<int>(i);
}

LinW Engine::lin(const std::string& wname, const std::string& bname, int K, int row0, int rows) {
  LinW l;
  TensorEntry* w = find(wname);
  const int N = rows < 0 ? static_cast<int>(w->shape[0]) : rows;
  l.N = N; l.K = K;
  l.w = shadow_ + w->offset + 1LL * row0 * K;
  l.w32 = params_ + w->offset + 1LL * row0 * K;
  l.gw = grads_ ? grads_ + w->offset + 1LL * row0 * K : nullptr;
  if (!bname.empty()) { TensorEntry* b = find(bname); l.b = params_ + b->offset + row0; l.gb = grads_ ? grads_ + b->offset + row0 : nullptr; }
  return l;
}
LnW Engine::lnw(const std::string& prefix) {
  LnW l;
  l.g = P(prefix + "weight"); l.b = P(prefix + "bias"); l.gg = G(prefix + "weight"); l.gb = G(prefix + "bias");
  return l;
}
ConvW Engine::convw(const std::string& wname, const std::string& bnp, int Cin, int Cout, int KH, int stride, int pad) {
  ConvW c;
  c.Cin = Cin; c.Cout = Cout; c.KH = KH; c.stride = stride; c.pad = pad;
  c.w32 = P(wname); c.gw = G(wname);
  c.bn[0] = P(bnp + "weight"); c.bn[1] = P(bnp + "bias"); c.bn[2] = P(bnp + "running_mean"); c.bn[3] = P(bnp + "running_var");
  return c;
}

void Engine::setup_views() {
  const int d = cfg_.hidden_dim, ff = cfg_.dim_ff;
  auto enc = [&](const std::string& p) {
    EncW w;
    w.sa.qk = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, 0, 2 * d);
    w.sa.q = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, 0, d);
    w.sa.k = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, d, d);
    w.sa.v = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, 2 * d, d);
    w.sa.o = lin(p + "self_attn.out_proj.weight", p + "self_attn.out_proj.bias", d);
    w.l1 = lin(p + "linear1.weight", p + "linear1.bias", d);
    w.l2 = lin(p + "linear2.weight", p + "linear2.bias", ff);
    w.n1 = lnw(p + "norm1."); w.n2 = lnw(p + "norm2.");
    return w;
  };
  enc_w_.clear(); cvae_w_.clear(); dec_w_.clear();
  for (int i = 0; i < cfg_.enc_layers; ++i) enc_w_.push_back(enc("transformer.encoder.layers." + std::to_string(i) + "."));
  if (!cfg_.no_encoder) for (int i = 0; i < cfg_.enc_layers; ++i) cvae_w_.push_back(enc("encoder.layers." + std::to_string(i) + "."));
  for (int i = 0; i < cfg_.dec_layers; ++i) {
    const std::string p = "transformer.decoder.layers." + std::to_string(i) + ".";
    DecW w;
    w.sa.qk = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, 0, 2 * d);
    w.sa.v = lin(p + "self_attn.in_proj_weight", p + "self_attn.in_proj_bias", d, 2 * d, d);
    w.sa.o = lin(p + "self_attn.out_proj.weight", p + "self_attn.out_proj.bias", d);
    w.ca.q = lin(p + "multihead_attn.in_proj_weight", p + "multihead_attn.in_proj_bias", d, 0, d);
    w.ca.k = lin(p + "multihead_attn.in_proj_weight", p + "multihead_attn.in_proj_bias", d, d, d);
    w.ca.v = lin(p + "multihead_attn.in_proj_weight", p + "multihead_attn.in_proj_bias", d, 2 * d, d);
    w.ca.o = lin(p + "multihead_attn.out_proj.weight", p + "multihead_attn.out_proj.bias", d);
    w.l1 = lin(p + "linear1.weight", p + "linear1.bias", d);
    w.l2 = lin(p + "linear2.weight", p + "linear2.bias", ff);
    w.n1 = lnw(p + "norm1."); w.n2 = lnw(p + "norm2."); w.n3 = lnw(p + "norm3.");
    dec_w_.push_back(w);
  }
  dec_norm_ = lnw("transformer.decoder.norm.");
  input_proj_ = lin("input_proj.weight", "input_proj.bias", 512);
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    const std::string p = "backbones." + std::to_string(k) + ".0.body.";
    CamW& c = cam_w_[k];
    ConvW keep = c.stem;
    c.stem = convw(p + "conv1.weight", p + "bn1.", 3, 64, 7, 2, 3);
    c.stem.wpk = keep.wpk; c.stem.scale = keep.scale; c.stem.shift = keep.shift;
    for (int s = 0; s < 4; ++s)
      for (int b = 0; b < 2; ++b) {
        BlockW& bw = c.blk[s * 2 + b];
        const std::string q = p + kStages[s] + "." + std::to_string(b) + ".";
        const int cin = b == 0 ? kStageCin[s] : kStageCout[s], cout = kStageCout[s], st = b == 0 ? kStageStride[s] : 1;
        auto carry = [&](ConvW& dst, ConvW src) {
          src.wpk = dst.wpk; src.wpkt = dst.wpkt; src.scale = dst.scale; src.shift = dst.shift;
          src.wpk2 = dst.wpk2; src.wpkt2 = dst.wpkt2; src.scale2 = dst.scale2; src.shift2 = dst.shift2;
          dst = src;
        };
        carry(bw.c1, convw(q + "conv1.weight", q + "bn1.", cin, cout, 3, st, 1));
        carry(bw.c2, convw(q + "conv2.weight", q + "bn2.", cout, cout, 3, 1, 1));
        bw.has_ds = (b == 0 && (st != 1 || cin != cout));
        if (bw.has_ds) carry(bw.ds, convw(q + "downsample.0.weight", q + "downsample.1.", cin, cout, 1, st, 0));
      }
  }
}

int Engine::bind(float* params, float* grads, float* exp_avg, float* exp_avg_sq) {
  ACT_CHECK(params != nullptr, "params must not be null");
  ACT_CHECK((reinterpret_cast<uintptr_t>(params) & 15) == 0, "flat parameter buffer must be 16-byte aligned");
  params_ = params; grads_ = grads; m_ = exp_avg; v_ = exp_avg_sq;
  setup_views();
  return 0;
}

int Engine::refresh_shadows(cudaStream_t st) {
  ACT_CHECK(params_ != nullptr, "engine not bound");
  // linear / embedding operands: flat bf16 mirror of the trainable regions (AdamW keeps it fresh afterwards)
  const long long n_mirror = region_off_[R_BUFFER];
  if (cast_bf16(params_, shadow_, n_mirror, st)) return -1;
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    CamW& c = cam_w_[k];
    if (bn_fold(c.stem.bn[0], c.stem.bn[1], c.stem.bn[2], c.stem.bn[3], c.stem.scale, c.stem.shift, 64, st)) return -1;
    // entries 64*j.. repeat 0..63: the grouped stem GEMM has N = 64*g (g output pixels per row)
    for (int j = 1; j < 4; ++j) {
      ACT_CUDA(cudaMemcpyAsync(c.stem.scale + 64 * j, c.stem.scale, 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
      ACT_CUDA(cudaMemcpyAsync(c.stem.shift + 64 * j, c.stem.shift, 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    if (pack_stem_weights(c.stem.w32, c.stem.wpk, st)) return -1;
    for (int i = 0; i < 8; ++i) {
      BlockW& b = c.blk[i];
      ConvW* cs[3] = {&b.c1, &b.c2, b.has_ds ? &b.ds : nullptr};
      for (ConvW* w : cs) {
        if (!w) continue;
        if (bn_fold(w->bn[0], w->bn[1], w->bn[2], w->bn[3], w->scale, w->shift, w->Cout, st)) return -1;
        if (pack_conv_weights(w->w32, w->scale, w->wpk, w->wpkt, w->Cout, w->Cin, w->KH * w->KH, st)) return -1;
        if (w->scale2) {
          for (int h = 0; h < 2; ++h) {
            ACT_CUDA(cudaMemcpyAsync(w->scale2 + 64 * h, w->scale, 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
            ACT_CUDA(cudaMemcpyAsync(w->shift2 + 64 * h, w->shift, 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
          }
        }
      }
    }
    if (repack_pair_operands(k, st)) return -1;
  }
  return 0;
}

// paired 64-channel operands (forward from wpk, data gradient from wpkt) of one camera, one launch
int Engine::repack_pair_operands(int cam, cudaStream_t st) {
  PairJobs jobs;
  for (int i = 0; i < 8; ++i) {
    BlockW& b = cam_w_[cam].blk[i];
    for (ConvW* w : {&b.c1, &b.c2}) {
      if (!w->wpk2) continue;
      for (int dg = 0; dg < (cfg_.train ? 2 : 1); ++dg) {
        PairJob& j = jobs.job[jobs.n++];
        j.src = dg ? w->wpkt : w->wpk; j.dst = dg ? w->wpkt2 : w->wpk2;
        conv_pair_slots(dg != 0, j.even_slot, j.odd_slot);
      }
    }
  }
  if (jobs.n == 0) return 0;
  ++launches_;
  return pack_pair_batched(jobs, st);
}

int Engine::zero_grad(cudaStream_t st, bool async) {
  ACT_CHECK(grads_ != nullptr, "no gradient buffer bound");
  const long long n = region_off_[R_DEAD];     // MAIN | CVAE | BACKBONE are contiguous and first
  // async: the clear runs on the weight-gradient stream beside the forward pass; backward()/adamw() join it
  cudaStream_t s = async ? wg_fork(st) : st;
  ACT_CUDA(cudaMemsetAsync(grads_, 0, n * sizeof(float), s));
  return 0;
}

int Engine::begin_opt_step(int step, cudaStream_t st) {
  ACT_CHECK(step >= 0, "optimizer step must be >= 0 (0 = advance the device counter)");
  ++launches_;
  return set_counter(opt_step_dev_, step, step == 0, st);
}

int Engine::adamw_range(long long off, long long n, float lr, bool shadow, int step, cudaStream_t st, int blocks_per_sm) {
  ACT_CHECK(params_ && grads_ && m_ && v_, "optimizer state not bound");
  if (n <= 0) return 0;
  AdamWArgs a;
  a.step_dev = opt_step_dev_; a.abort_flag = dp_status_;
  a.p = params_ + off; a.g = grads_ + off; a.m = m_ + off; a.v = v_ + off; a.shadow = shadow ? shadow_ + off : nullptr;
  a.n = n; a.lr = lr; a.beta1 = cfg_.beta1; a.beta2 = cfg_.beta2; a.eps = cfg_.adam_eps; a.wd = cfg_.weight_decay;
  a.step = step; a.mode = 0; a.background = blocks_per_sm;
  ++launches_;
  if (dp_fused_) {                                          // reduce + update + broadcast over peer memory
    dp_ranges_.emplace_back(off, n);
    if (dp_defer_now_) { dp_deferred_.push_back(DpSeg{a, off, st, false}); return 0; }
    return dp_adamw_step(peers_, a, off, dp_stream_);
  }
  return adamw_step(a, st);
}
int Engine::adamw_region(int r, int step, cudaStream_t st, int blocks_per_sm) {
  ACT_CHECK(params_ && grads_ && m_ && v_, "optimizer state not bound");
  AdamWArgs a;
  a.step_dev = opt_step_dev_; a.abort_flag = dp_status_;
  const long long o = region_off_[r];
  const bool shadow = r == R_MAIN || r == R_CVAE;
  a.p = params_ + o; a.g = grads_ + o; a.m = m_ + o; a.v = v_ + o; a.shadow = shadow ? shadow_ + o : nullptr;
  a.n = region_n_[r]; a.lr = r == R_BACKBONE ? cfg_.lr_backbone : cfg_.lr; a.beta1 = cfg_.beta1; a.beta2 = cfg_.beta2;
  a.eps = cfg_.adam_eps; a.wd = cfg_.weight_decay; a.step = step; a.mode = r == R_DEAD ? 1 : 0;
  a.background = blocks_per_sm;
  ++launches_;
  if (dp_fused_ && r != R_DEAD) {
    dp_ranges_.emplace_back(o, a.n);
    if (dp_defer_now_) {
      // (the conv re-pack may only be armed once this range's kernel is on the exchange stream: dp_flush_deferred)
      dp_deferred_.push_back(DpSeg{a, o, st, r == R_BACKBONE});
      return 0;
    }
    if (dp_adamw_step(peers_, a, o, dp_stream_)) return -1;
    // the conv operands are re-packed from the fp32 masters once every rank's slice of them has landed here: that is
    // what the step's closing barrier guarantees, so the re-pack is deferred to it (repack_conv_operands)
    if (r == R_BACKBONE) dp_pack_pending_ = true;
    return 0;
  }
  if (adamw_step(a, st)) return -1;
  if (r != R_BACKBONE) return 0;
  return repack_conv_operands(st);
}

int Engine::repack_conv_operands(cudaStream_t st, int cam, bool stem) {
  // packed conv operands (transposed / scaled layouts) are rebuilt from the fp32 masters, one launch per camera
  for (int k = 0; k < cfg_.num_cameras; ++k) {
    if (cam >= 0 && k != cam) continue;
    CamW& c = cam_w_[k];
    if (stem) {
      if (pack_stem_weights(c.stem.w32, c.stem.wpk, st)) return -1;
      ++launches_;
    }
    PackJobs jobs;
    for (int i = 0; i < 8; ++i) {
      BlockW& b = c.blk[i];
      ConvW* cs[3] = {&b.c1, &b.c2, b.has_ds ? &b.ds : nullptr};
      for (ConvW* w : cs) {
        if (!w) continue;
        jobs.job[jobs.n++] = PackJob{w->w32, w->scale, w->wpk, w->wpkt, w->Cout, w->Cin, w->KH * w->KH, 0};
      }
    }
    if (pack_conv_batched(jobs, st)) return -1;
    ++launches_;
    if (repack_pair_operands(k, st)) return -1;
  }
  return 0;
}

// ------------------------------------------------------------------------------------------ peer-memory data parallel
int Engine::dp_local(void** grads, void** params, void** shadow, void** signal) {
  ACT_CHECK(params_ && grads_ && shadow_, "dp: parameters / gradients not bound");
  if (!dp_signal_) dp_signal_ = reinterpret_cast<uint32_t*>(alloc<float>(DP_MAX_SITES * DP_MAX_WORLD));      // own cudaMalloc, zeroed
  if (!dp_epoch_) {
    dp_epoch_ = reinterpret_cast<uint32_t*>(alloc<float>(DP_MAX_SITES));
    dp_status_ = reinterpret_cast<uint32_t*>(alloc<float>(4));
    ACT_CHECK(dp_signal_ && dp_epoch_ && dp_status_, "dp: flag allocation failed");
    ACT_CUDA(cudaDeviceSynchronize());                                                          // alloc's memset
  }
  *grads = grads_; *params = params_; *shadow = shadow_; *signal = dp_signal_;
  return 0;
}

int Engine::dp_use_buffers(void* shadow, void* signal) {
  ACT_CHECK(params_ && shadow_, "dp_use_buffers: engine not bound");
  ACT_CHECK(peers_.world <= 1, "dp_use_buffers: call before the peers are connected");
  ACT_CHECK(dp_signal_ == nullptr || signal == nullptr || signal == dp_signal_, "dp_use_buffers: flag block already allocated");
  ACT_CHECK((reinterpret_cast<uintptr_t>(shadow) & 255) == 0 && (reinterpret_cast<uintptr_t>(signal) & 15) == 0, "dp_use_buffers: alignment");
  ACT_CUDA(cudaDeviceSynchronize());
  if (shadow && shadow != shadow_) {
    ACT_CUDA(cudaMemcpy(shadow, shadow_, sizeof(bf16) * total_floats_, cudaMemcpyDeviceToDevice));
    shadow_ = static_cast<bf16*>(shadow);                      // the workspace copy stays allocated but unused
    setup_views();
  }
  if (signal && !dp_signal_) {
    void *g, *p, *s, *sg;
    dp_signal_ = static_cast<uint32_t*>(signal);
    ACT_CUDA(cudaMemset(dp_signal_, 0, DP_SIGNAL_BYTES));
    if (dp_local(&g, &p, &s, &sg)) return -1;                  // allocates the (local) epoch / status words
  }
  return 0;
}

int Engine::dp_set_multicast(void* mc_grads, void* mc_params, void* mc_shadow) {
  ACT_CHECK(peers_.world > 1, "dp_set_multicast: connect the peers first");
  ACT_CHECK(((reinterpret_cast<uintptr_t>(mc_grads) | reinterpret_cast<uintptr_t>(mc_params) | reinterpret_cast<uintptr_t>(mc_shadow)) & 15) == 0,
            "dp_set_multicast: multicast addresses must be 16-byte aligned");
  ACT_CHECK((mc_grads != nullptr) == (mc_params != nullptr) && (mc_grads != nullptr) == (mc_shadow != nullptr),
            "dp_set_multicast: give all three multicast addresses or none");
  peers_.mc_grads = static_cast<float*>(mc_grads); peers_.mc_params = static_cast<float*>(mc_params);
  peers_.mc_shadow = static_cast<bf16*>(mc_shadow);
  return 0;
}

namespace {
struct DpBlob { cudaIpcMemHandle_t h[4]; long long off[4]; int device; int pad; };
static_assert(sizeof(DpBlob) == Engine::DP_BLOB_BYTES, "DpBlob layout");

// base address of the allocation that contains p (cudaIpcGetMemHandle wants the base; torch sub-allocates)
int alloc_base(const void* p, void** base) {
  using Fn = CUresult (*)(CUdeviceptr*, size_t*, CUdeviceptr);
  static Fn fn = nullptr;
  if (!fn) {
    void* q = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &q, cudaEnableDefault, &qr) != cudaSuccess || !q) {
      set_error("cuMemGetAddressRange unavailable"); return -1;
    }
    fn = reinterpret_cast<Fn>(q);
  }
  CUdeviceptr b = 0; size_t sz = 0;
  if (fn(&b, &sz, reinterpret_cast<CUdeviceptr>(p)) != CUDA_SUCCESS) { set_error("cuMemGetAddressRange failed"); return -1; }
  *base = reinterpret_cast<void*>(b);
  return 0;
}
}  // namespace

int Engine::dp_export(void* blob, int cap) {
  ACT_CHECK(cap >= DP_BLOB_BYTES, "dp_export: buffer too small");
  void* ptr[4];
  if (dp_local(&ptr[0], &ptr[1], &ptr[2], &ptr[3])) return -1;
  DpBlob b;
  memset(&b, 0, sizeof b);
  ACT_CUDA(cudaGetDevice(&b.device));
  for (int i = 0; i < 4; ++i) {
    void* base = nullptr;
    if (alloc_base(ptr[i], &base)) return -1;
    ACT_CUDA(cudaIpcGetMemHandle(&b.h[i], base));
    b.off[i] = static_cast<const char*>(ptr[i]) - static_cast<const char*>(base);
  }
  memcpy(blob, &b, sizeof b);
  return DP_BLOB_BYTES;
}

int Engine::dp_connect(int rank, int world, const void* blobs, int blob_bytes) {
  ACT_CHECK(blob_bytes == DP_BLOB_BYTES, "dp_connect: blob size mismatch");
  ACT_CHECK(world >= 2 && world <= DP_MAX_WORLD && rank >= 0 && rank < world, "dp_connect: bad rank / world");
  void* mine[4];
  if (dp_local(&mine[0], &mine[1], &mine[2], &mine[3])) return -1;
  void* tab[4][DP_MAX_WORLD] = {};
  for (int j = 0; j < world; ++j) {
    if (j == rank) { for (int i = 0; i < 4; ++i) tab[i][j] = mine[i]; continue; }
    DpBlob b;
    memcpy(&b, static_cast<const char*>(blobs) + 1LL * j * DP_BLOB_BYTES, sizeof b);
    for (int i = 0; i < 4; ++i) {
      void* base = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&base, b.h[i], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        set_error(std::string("cudaIpcOpenMemHandle (rank ") + std::to_string(j) + "): " + cudaGetErrorString(e));
        cudaGetLastError();
        return -1;
      }
      dp_opened_.push_back(base);
      tab[i][j] = static_cast<char*>(base) + b.off[i];
    }
  }
  return dp_set_peers(rank, world, tab[0], tab[1], tab[2], tab[3]);
}

int Engine::dp_set_peers(int rank, int world, void* const* grads, void* const* params, void* const* shadow, void* const* signal) {
  ACT_CHECK(world >= 2 && world <= DP_MAX_WORLD && rank >= 0 && rank < world, "dp_set_peers: bad rank / world");
  void* mine[4];
  if (dp_local(&mine[0], &mine[1], &mine[2], &mine[3])) return -1;
  ACT_CHECK(grads[rank] == mine[0] && params[rank] == mine[1] && shadow[rank] == mine[2] && signal[rank] == mine[3],
            "dp_set_peers: entry `rank` must be this engine's own buffers");
  peers_ = DpPeers();
  peers_.rank = rank; peers_.world = world;
  for (int j = 0; j < world; ++j) {
    ACT_CHECK(grads[j] && params[j] && shadow[j] && signal[j], "dp_set_peers: null peer pointer");
    peers_.grads[j] = static_cast<float*>(grads[j]); peers_.params[j] = static_cast<float*>(params[j]);
    peers_.shadow[j] = static_cast<bf16*>(shadow[j]); peers_.signal[j] = static_cast<uint32_t*>(signal[j]);
  }
  if (!dp_stream_) {
    ACT_CUDA(cudaStreamCreateWithFlags(&dp_stream_, cudaStreamNonBlocking));
    ACT_CUDA(cudaStreamCreateWithFlags(&dp_pack_stream_, cudaStreamNonBlocking));
    ACT_CUDA(cudaEventCreateWithFlags(&dp_pack_ev_, cudaEventDisableTiming));
    ACT_CUDA(cudaEventCreateWithFlags(&dp_pack_done_, cudaEventDisableTiming));
    for (cudaEvent_t& e : dp_ev_) ACT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    ACT_CUDA(cudaEventCreateWithFlags(&dp_done_, cudaEventDisableTiming));
  }
  return 0;
}

// One barrier for every deferred range (each rank's producers of all of them are waited for first), then their exchange
// kernels side by side on the exchange stream and its helpers: the link sees one long transfer instead of three short
// ones with a barrier and a ramp between them.
int Engine::dp_flush_deferred() {
  if (dp_deferred_.empty()) return 0;
  ACT_CHECK(dp_deferred_.size() <= 4, "dp: too many deferred ranges");
  if (!dp_aux_[0]) {
    for (cudaStream_t& s : dp_aux_) ACT_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    for (cudaEvent_t& e : dp_aux_ev_) ACT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  for (size_t k = 1; k < dp_deferred_.size(); ++k) {
    if (dp_deferred_[k].producer == dp_deferred_[0].producer || dp_deferred_[k].producer == dp_stream_) continue;
    ACT_CUDA(cudaEventRecord(dp_aux_ev_[k], dp_deferred_[k].producer));
    ACT_CUDA(cudaStreamWaitEvent(dp_stream_, dp_aux_ev_[k], 0));
  }
  if (dp_sync_point(dp_deferred_[0].producer)) return -1;
  ACT_CUDA(cudaEventRecord(dp_aux_ev_[0], dp_stream_));
  for (size_t k = 0; k < dp_deferred_.size(); ++k) {
    cudaStream_t s = k == 0 ? dp_stream_ : dp_aux_[k - 1];
    if (k > 0) ACT_CUDA(cudaStreamWaitEvent(s, dp_aux_ev_[0], 0));
    ++launches_;
    if (dp_adamw_step(peers_, dp_deferred_[k].a, dp_deferred_[k].off, s)) return -1;
    if (dp_deferred_[k].repack) dp_pack_pending_ = true;     // re-pack behind the NEXT barrier (every rank's slice has landed)
    if (k > 0) {
      ACT_CUDA(cudaEventRecord(dp_aux_ev_[4 + k], s));
      ACT_CUDA(cudaStreamWaitEvent(dp_stream_, dp_aux_ev_[4 + k], 0));
    }
  }
  dp_deferred_.clear();
  return 0;
}

int Engine::dp_sync_point(cudaStream_t producer) {
  ACT_CHECK(peers_.world > 1 && dp_stream_, "dp: peers not connected");
  ACT_CHECK(dp_site_ < DP_MAX_SITES, "dp: too many barrier sites in one step");
  if (producer != dp_stream_) {
    ACT_CUDA(cudaEventRecord(dp_ev_[dp_site_], producer));
    ACT_CUDA(cudaStreamWaitEvent(dp_stream_, dp_ev_[dp_site_], 0));
  }
  ++launches_;
  if (dp_barrier(peers_, dp_epoch_, dp_status_, dp_site_++, dp_stream_)) return -1;
  if (dp_pack_pending_) {
    // every rank has passed this barrier, so every rank's backbone kernel (earlier on its exchange stream) has finished
    // storing its slice here: rebuild the packed conv operands on a stream of their own, beside the remaining exchange
    dp_pack_pending_ = false;
    ACT_CUDA(cudaEventRecord(dp_pack_ev_, dp_stream_));
    ACT_CUDA(cudaStreamWaitEvent(dp_pack_stream_, dp_pack_ev_, 0));
    if (repack_conv_operands(dp_pack_stream_)) return -1;
    ACT_CUDA(cudaEventRecord(dp_pack_done_, dp_pack_stream_));
    dp_pack_running_ = true;
  }
  return 0;
}

int Engine::dp_preset_signal(int peer_rank, uint32_t epoch) {
  ACT_CHECK(dp_signal_ != nullptr && peer_rank >= 0 && peer_rank < DP_MAX_WORLD, "dp_preset_signal: no flag block / bad rank");
  std::vector<uint32_t> h(DP_MAX_SITES * DP_MAX_WORLD);
  ACT_CUDA(cudaMemcpy(h.data(), dp_signal_, h.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  for (int s = 0; s < DP_MAX_SITES; ++s) h[s * DP_MAX_WORLD + peer_rank] = epoch;
  ACT_CUDA(cudaMemcpy(dp_signal_, h.data(), h.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  return 0;
}

int Engine::dp_status() {
  if (!dp_status_) return 0;
  uint32_t s = 0;
  ACT_CUDA(cudaMemcpy(&s, dp_status_, sizeof s, cudaMemcpyDeviceToHost));
  if (s) set_error("data-parallel flag barrier timed out (a peer did not reach the same step)");
  return static_cast<int>(s);
}

int Engine::adamw(int step, cudaStream_t st, bool begin) {
  if (wg_join(st)) return -1;
  if (begin && begin_opt_step(step, st)) return -1;
  const int order[4] = {R_MAIN, R_CVAE, R_BACKBONE, R_DEAD};
  for (int r : order)
    if (adamw_region(r, step, st)) return -1;
  return 0;
}

}  // namespace act
