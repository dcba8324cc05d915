// Convolution forward / data-gradient / weight-gradient as tap tables over the shared tcgen05 kernel.
// Activations NHWC bf16; packed weights:
//   forward  : wpk  [Cout][(kh*KW+kw)*Cin + ci]
//   dgrad    : wpkt [Cin ][(kh*KW+kw)*Cout + co]   (optionally pre-multiplied by the FrozenBN scale of co)
// Stride-2 access goes through four parity planes of the NHWC tensor (plain tiled TMA maps, zero OOB fill = padding).
#include "conv_ops.h"

namespace act {

static int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
static int pmod(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }

static void full_plane(PlaneDesc* pl, const bf16* ptr, int N, int H, int W, int C) {
  pl->ptr = ptr; pl->C = C; pl->W = W; pl->H = H; pl->N = N;
  pl->sW = C; pl->sH = 1LL * W * C; pl->sN = 1LL * H * W * C;
}
// parity plane (ph,pw) of an [N,H,W,C] tensor with even H, W
static void parity_plane(PlaneDesc* pl, const bf16* ptr, int N, int H, int W, int C, int ph, int pw) {
  pl->ptr = ptr + (1LL * ph * W + pw) * C; pl->C = C; pl->W = W / 2; pl->H = H / 2; pl->N = N;
  pl->sW = 2LL * C; pl->sH = 2LL * W * C; pl->sN = 1LL * H * W * C;
}

int conv_out_dim(int in, int k, int stride, int pad) { return (in + 2 * pad - k) / stride + 1; }

// Builds planes + taps for reading the conv *input* x at the positions each output pixel needs.
static int input_taps(const ConvShape& s, const bf16* x, PlaneDesc* planes, int* nplanes, TapDesc* taps, int* ntaps) {
  ACT_CHECK(s.stride == 1 || s.stride == 2, "stride must be 1 or 2");
  ACT_CHECK(s.KH == 1 || s.KH == 3, "kernel must be 1x1 or 3x3");
  if (s.stride == 1) {
    full_plane(&planes[0], x, s.N, s.H, s.W, s.Cin);
    *nplanes = 1;
  } else {
    ACT_CHECK(s.H % 2 == 0 && s.W % 2 == 0, "stride-2 conv needs even input dims");
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw) parity_plane(&planes[ph * 2 + pw], x, s.N, s.H, s.W, s.Cin, ph, pw);
    *nplanes = 4;
  }
  int n = 0;
  for (int kh = 0; kh < s.KH; ++kh)
    for (int kw = 0; kw < s.KH; ++kw) {
      TapDesc t;
      const int oy = kh - s.pad, ox = kw - s.pad;
      if (s.stride == 1) { t.map = 0; t.dh = oy; t.dw = ox; }
      else { t.map = pmod(oy, 2) * 2 + pmod(ox, 2); t.dh = floordiv(oy, 2); t.dw = floordiv(ox, 2); }
      t.wslot = kh * s.KH + kw;
      taps[n++] = t;
    }
  *ntaps = n;
  return 0;
}

// ---- two output pixels per GEMM row for 64 -> 64 channel 3x3/s1 convs (forward and data gradient).
// An M=128 tcgen05.mma costs the same for N = 64 and N = 128, so these convs (ResNet layer1: the largest images of the
// trunk) ran at half the tensor rate.  View the NHWC-64 tensors as pixel PAIRS, [N,H,W/2,128]: the conv becomes a dense
// 3x3/s1 conv with 128 -> 128 channels over pairs (so it runs on the halo-resident conv3_kernel unchanged) whose weight
// has structurally zero blocks: the pair to the left contributes only its odd pixel, the pair to the right only its even
// pixel.  Of the 9 taps x 2 channel blocks, 6 are zero and are skipped (ConvLaunch::skip_mask): 48 N=128 MMAs per 256
// output pixels instead of 72 N=64 MMAs.
bool conv_pair_ok(const ConvShape& s) {
  static int env = -1;
  if (env < 0) { const char* e = getenv("ACT_CONV_PAIR"); env = (e && e[0] == '0') ? 0 : 1; }
  return env && s.KH == 3 && s.stride == 1 && s.pad == 1 && s.Cin == 64 && s.Cout == 64 && s.W % 2 == 0;
}
// forward tap (kh,kw) reads column offset dw = kw - 1, row offset dh = kh - 1; dgrad tap reads dw = 1 - kw, dh = 1 - kh.
// Block sp = ((dh+1)*3 + (dwp+1))*2 + pi holds input pixel e = 2*dwp + pi relative to the pair's even pixel.
void conv_pair_slots(bool dgrad, int8_t even_slot[18], int8_t odd_slot[18]) {
  for (int dh = -1; dh <= 1; ++dh)
    for (int dwp = -1; dwp <= 1; ++dwp)
      for (int pi = 0; pi < 2; ++pi) {
        const int sp = ((dh + 1) * 3 + (dwp + 1)) * 2 + pi, e = 2 * dwp + pi;
        auto slot = [&](int dw) -> int8_t {
          if (dw < -1 || dw > 1) return -1;
          const int kh = dgrad ? 1 - dh : dh + 1, kw = dgrad ? 1 - dw : dw + 1;
          return static_cast<int8_t>(kh * 3 + kw);
        };
        even_slot[sp] = slot(e);
        odd_slot[sp] = slot(e - 1);
      }
}
static int pair_launch(const bf16* in, int N, int H, int W, const bf16* wpair, const EpiDesc& epi, cudaStream_t stream) {
  ConvLaunch c;
  full_plane(&c.planes[0], in, N, H, W / 2, 128);
  c.nplanes = 1;
  int n = 0;
  for (int dh = -1; dh <= 1; ++dh)
    for (int dwp = -1; dwp <= 1; ++dwp) {
      TapDesc t;
      t.map = 0; t.dh = static_cast<int16_t>(dh); t.dw = static_cast<int16_t>(dwp); t.wslot = static_cast<int16_t>(n);
      // left pair: only its odd pixel (channel block 1) is within reach; right pair: only its even pixel (block 0)
      if (dwp == -1) c.skip_mask |= 1u << (n * 2 + 0);
      if (dwp == +1) c.skip_mask |= 1u << (n * 2 + 1);
      c.taps[n++] = t;
    }
  c.ntaps = n;
  c.Cin = 128; c.Cout = 128; c.wgt = wpair; c.wgt_ld = 9 * 128;
  c.vW = W / 2; c.vH = H; c.NIMG = N;
  c.og = PixGeom{H, W / 2, 1, 1, 0, 0, 128};
  c.rg = c.og;
  c.epi = epi;
  c.algo_flops = 2.0 * N * H * W * 64 * 64 * 9;
  return tc_conv(c, stream);
}

int conv_fwd(const ConvShape& s, const bf16* x, const bf16* wpk, const EpiDesc& epi, cudaStream_t stream, const bf16* wpair) {
  if (wpair && conv_pair_ok(s)) return pair_launch(x, s.N, s.H, s.W, wpair, epi, stream);
  ConvLaunch c;
  if (input_taps(s, x, c.planes, &c.nplanes, c.taps, &c.ntaps)) return -1;
  const int Ho = conv_out_dim(s.H, s.KH, s.stride, s.pad), Wo = conv_out_dim(s.W, s.KH, s.stride, s.pad);
  c.Cin = s.Cin; c.Cout = s.Cout; c.wgt = wpk; c.wgt_ld = 1LL * s.KH * s.KH * s.Cin;
  c.vW = Wo; c.vH = Ho; c.NIMG = s.N;
  c.og = PixGeom{Ho, Wo, 1, 1, 0, 0, s.Cout};
  c.rg = c.og;
  c.epi = epi;
  return tc_conv(c, stream);
}

// dx[n,y,x,ci] = sum_{kh,kw,co} dy[n,(y+pad-kh)/s,(x+pad-kw)/s,co] * W[co,ci,kh,kw]   (exact divisions only)
// epi.res_* (optional) is indexed like dx when stride == 1; for stride 2 `res_compact` is a [N,H/2,W/2,Cin]
// tensor added to parity class (0,0) only (the 1x1/s2 downsample branch).  epi.aux_bf16 masks with dx geometry.
int conv_dgrad(const ConvShape& s, const bf16* dy, const bf16* wpkt, EpiDesc epi, const bf16* res_compact,
               cudaStream_t stream, const bf16* wpair) {
  if (wpair && conv_pair_ok(s) && !res_compact) return pair_launch(dy, s.N, s.H, s.W, wpair, epi, stream);
  const int Ho = conv_out_dim(s.H, s.KH, s.stride, s.pad), Wo = conv_out_dim(s.W, s.KH, s.stride, s.pad);
  const int nclass = (s.stride == 1) ? 1 : 4;
  for (int cls = 0; cls < nclass; ++cls) {
    const int py = cls >> 1, px = cls & 1;
    ConvLaunch c;
    full_plane(&c.planes[0], dy, s.N, Ho, Wo, s.Cout);
    c.nplanes = 1;
    int n = 0;
    for (int kh = 0; kh < s.KH; ++kh)
      for (int kw = 0; kw < s.KH; ++kw) {
        const int ay = py + s.pad - kh, ax = px + s.pad - kw;
        if (s.stride == 2 && ((ay & 1) || (ax & 1))) continue;
        TapDesc t;
        t.map = 0; t.dh = (s.stride == 1) ? ay : ay / 2; t.dw = (s.stride == 1) ? ax : ax / 2;   // ay even -> exact
        t.wslot = kh * s.KH + kw;
        c.taps[n++] = t;
      }
    c.ntaps = n;
    c.Cin = s.Cout; c.Cout = s.Cin; c.wgt = wpkt; c.wgt_ld = 1LL * s.KH * s.KH * s.Cout;
    c.vW = s.W / s.stride; c.vH = s.H / s.stride; c.NIMG = s.N;
    c.og = PixGeom{s.H, s.W, s.stride, s.stride, py, px, s.Cin};
    c.rg = c.og;
    c.epi = epi;
    if (s.stride == 2) {
      c.epi.flags &= ~(EPI_RES_F32 | EPI_RES_BF16);
      c.epi.res_bf16 = nullptr; c.epi.res_f32 = nullptr;
      if (cls == 0 && res_compact) {
        c.epi.flags |= EPI_RES_BF16; c.epi.res_bf16 = res_compact;
        c.rg = PixGeom{s.H / 2, s.W / 2, 1, 1, 0, 0, s.Cin};
      }
    }
    if (n == 0) {
      // no tap reaches this parity class (1x1/s2): dx there is the residual or zero -> handled by caller's memset
      continue;
    }
    if (tc_conv(c, stream)) return -1;
  }
  return 0;
}

// ---- weight gradient: tc_wgrad accumulates into a staging buffer T[KK][Cout][Cin] (each warp-level reduction covers one
// contiguous 128-byte line instead of nine lines at stride KK); this pass folds it into dW.
static float* g_stage = nullptr;
static long long g_stage_n = 0;
int conv_wgrad_reserve(long long elems) {
  if (elems <= g_stage_n) return 0;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  (void)cs;
  if (g_stage) ACT_CUDA(cudaFree(g_stage));
  g_stage = nullptr; g_stage_n = 0;
  ACT_CUDA(cudaMalloc(&g_stage, sizeof(float) * elems));
  ACT_CUDA(cudaMemset(g_stage, 0, sizeof(float) * elems));
  g_stage_n = elems;
  return 0;
}

// block = 32(co) x 32(ci) x KK tile of one job: T rows along ci read coalesced, dW written in contiguous ci*KK runs.
// KK is a template constant (0 = runtime value) so the index arithmetic is shifts and constant multiplies: with a
// runtime divisor the kernel was instruction-bound (68 us for 11 M weights), not bandwidth-bound.
template <int KKT>
__device__ __forceinline__ void wgrad_fold_body(const FoldJob& jb, int b, float (*tile)[32 * 9 + 1]) {
  const int Cout = jb.Cout, Cin = jb.Cin, KK = KKT ? KKT : jb.KK;
  const int ci_tiles = (Cin + 31) / 32;
  const int co0 = (b / ci_tiles) * 32, ci0 = (b % ci_tiles) * 32;
  const int run = 32 * KK;
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const bool ci_ok = ci0 + lane < Cin;
  for (int t = 0; t < KK; ++t) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int co = wrp + 8 * c;
      float v = 0.f;
      if (ci_ok && co0 + co < Cout) v = __ldg(jb.T + (1LL * t * Cout + co0 + co) * Cin + ci0 + lane);
      tile[co][lane * KK + t] = v;
    }
  }
  __syncthreads();
  const int valid = min(32, Cin - ci0) * KK;                  // floats of a co row that exist in dW
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int co = wrp + 8 * c;
    if (co0 + co >= Cout) continue;
    const float a = jb.alpha * (jb.scale ? jb.scale[co0 + co] : 1.f);
    float* drow = jb.dw + (1LL * (co0 + co) * Cin + ci0) * KK;
    for (int o = lane; o < run; o += 32)
      if (o < valid) drow[o] += a * tile[co][o];
  }
}
__global__ void __launch_bounds__(256) wgrad_fold_kernel(const __grid_constant__ FoldJobs jobs) {
  pdl_trigger(); pdl_wait();
  __shared__ float tile[32][32 * 9 + 1];
  int b = blockIdx.x, j = 0;
  while (j + 1 < jobs.n && b >= jobs.job[j + 1].tile0) ++j;
  const FoldJob& jb = jobs.job[j];
  b -= jb.tile0;
  if (jb.KK == 9) wgrad_fold_body<9>(jb, b, tile);
  else if (jb.KK == 1) wgrad_fold_body<1>(jb, b, tile);
  else wgrad_fold_body<0>(jb, b, tile);
}
int wgrad_fold_batched(FoldJobs& jobs, cudaStream_t stream) {
  ACT_CHECK(jobs.n >= 1 && jobs.n <= FoldJobs::MAX, "wgrad_fold_batched: 1..64 jobs");
  int tiles = 0;
  for (int i = 0; i < jobs.n; ++i) {
    ACT_CHECK(jobs.job[i].KK >= 1 && jobs.job[i].KK <= 9, "wgrad_fold_batched: KK <= 9");
    jobs.job[i].tile0 = tiles;
    tiles += ((jobs.job[i].Cout + 31) / 32) * ((jobs.job[i].Cin + 31) / 32);
  }
  ACT_CUDA(launch_pdl(wgrad_fold_kernel, dim3(tiles), dim3(256), 0, stream, jobs));
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// staged weight gradient only: T[KK][Cout][Cin] += sum_pix dy[pix,co] * x[pix(+tap),ci]  (T zeroed by the caller)
int conv_wgrad_staged(const ConvShape& s, const bf16* x, const bf16* dy, float* T, cudaStream_t stream) {
  WgradLaunch w;
  if (input_taps(s, x, w.planes, &w.nplanes, w.taps, &w.ntaps)) return -1;
  const int Ho = conv_out_dim(s.H, s.KH, s.stride, s.pad), Wo = conv_out_dim(s.W, s.KH, s.stride, s.pad);
  full_plane(&w.dy, dy, s.N, Ho, Wo, s.Cout);
  w.Cin = s.Cin; w.Cout = s.Cout; w.KK = s.KH * s.KH; w.dw = T; w.alpha = 1.f; w.scale = nullptr;
  return tc_wgrad(w, stream);
}

// dW[co,ci,kh,kw] += alpha * scale[co] * sum_pix dy[pix,co] * x[pix(+tap),ci]
int conv_wgrad(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
               cudaStream_t stream) {
  WgradLaunch w;
  if (input_taps(s, x, w.planes, &w.nplanes, w.taps, &w.ntaps)) return -1;
  const int Ho = conv_out_dim(s.H, s.KH, s.stride, s.pad), Wo = conv_out_dim(s.W, s.KH, s.stride, s.pad);
  full_plane(&w.dy, dy, s.N, Ho, Wo, s.Cout);
  const int KK = s.KH * s.KH;
  const long long n = 1LL * KK * s.Cin * s.Cout;
  if (n > g_stage_n) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    ACT_CHECK(cs == cudaStreamCaptureStatusNone, "conv_wgrad staging buffer must be reserved before stream capture");
    if (conv_wgrad_reserve(n)) return -1;
  }
  (void)w; (void)Ho; (void)Wo;
  ACT_CUDA(cudaMemsetAsync(g_stage, 0, sizeof(float) * n, stream));
  if (conv_wgrad_staged(s, x, dy, g_stage, stream)) return -1;
  FoldJobs jobs;
  jobs.n = 1; jobs.job[0] = FoldJob{g_stage, scale, dw, alpha, s.Cout, s.Cin, KK, 0};
  return wgrad_fold_batched(jobs, stream);
}

// ---- the same weight gradient for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.  The staging tile lives in a
// file-scope device array of a fixed, modest size (2.25 MiB, held by every GPU that loads the library until exit): it covers
// KH*KW*Cin*Cout <= 9*256*256, i.e. every resnet18 conv up to and including layer3.  Every call shares the array and
// re-zeroes the part it uses on `stream` first, so consecutive calls on one stream are independent; calls that overlap on the
// GPU (different streams, no ordering) race on it.
static __device__ __align__(256) float g_wgrad_static[CONV_WGRAD_NOALLOC_MAX];
static float* g_wgrad_static_ptr[64] = {};          // per device: the array's DEVICE address (its name is only a host stand-in)

int conv_wgrad_noalloc(const ConvShape& s, const bf16* x, const bf16* dy, float* dw, float alpha, const float* scale,
                       cudaStream_t stream) {
  const int KK = s.KH * s.KH;
  const long long n = 1LL * KK * s.Cin * s.Cout;
  ACT_CHECK(n >= 1 && n <= CONV_WGRAD_NOALLOC_MAX,
            "act_conv_wgrad_noalloc: KH*KW*Cin*Cout exceeds the static scratch (589824 floats = 9*256*256); use act_conv_wgrad");
  int dev = -1;
  ACT_CUDA(cudaGetDevice(&dev));
  ACT_CHECK(dev >= 0 && dev < 64, "act_conv_wgrad_noalloc: device ordinal out of range");
  if (!g_wgrad_static_ptr[dev]) {                   // set-up, once per device, with that device current
    float* ptr = nullptr;
    ACT_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_wgrad_static));
    ACT_CHECK(ptr != nullptr, "act_conv_wgrad_noalloc: cudaGetSymbolAddress returned a null scratch address");
    g_wgrad_static_ptr[dev] = ptr;
  }
  float* T = g_wgrad_static_ptr[dev];
  ACT_CUDA(cudaMemsetAsync(T, 0, sizeof(float) * n, stream));      // the array keeps the previous call's sums
  if (conv_wgrad_staged(s, x, dy, T, stream)) return -1;
  FoldJobs jobs;
  jobs.n = 1; jobs.job[0] = FoldJob{T, scale, dw, alpha, s.Cout, s.Cin, KK, 0};
  return wgrad_fold_batched(jobs, stream);
}

}  // namespace act
