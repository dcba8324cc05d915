// Host launchers for tc_kernel: tensor-map construction (driver entry point fetched at run time so the
// library loads on machines without libcuda), tile selection, launch.
#include "tc_api.h"
#include "conv3_kernel.cuh"
#include "attn_kernel.cuh"
#include <mutex>
#include <cstring>
#include <algorithm>
#include <vector>
#include <cstdio>
#include <cstdlib>

namespace act {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const char* last_error() { return g_err.c_str(); }

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("ACT_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
  return on == 1;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// bf16 tensor map, 4-D, 128B swizzle, zero OOB fill.  dims/box innermost first; strides (elements) for dims 1..3.
static int make_tmap(CUtensorMap* m, const bf16* ptr, const long long dims[4], const long long strides_el[3],
                     const int box[4]) {
  EncodeTiledFn enc = get_encode();
  ACT_CHECK(enc != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  cuuint64_t gd[4], gs[3];
  cuuint32_t bx[4], es[4] = {1, 1, 1, 1};
  for (int i = 0; i < 4; ++i) { gd[i] = static_cast<cuuint64_t>(std::max(1LL, dims[i])); bx[i] = static_cast<cuuint32_t>(box[i]); }
  for (int i = 0; i < 3; ++i) gs[i] = static_cast<cuuint64_t>(strides_el[i]) * 2;
  ACT_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "TMA base must be 16-byte aligned");
  for (int i = 0; i < 3; ++i) ACT_CHECK(gs[i] % 16 == 0 && gs[i] > 0, "TMA stride must be a positive multiple of 16 bytes");
  ACT_CHECK(box[0] * 2 == 128, "inner box must span the 128B swizzle");
  for (int i = 0; i < 4; ++i) ACT_CHECK(box[i] >= 1 && box[i] <= 256, "box dim out of range");
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16*>(ptr), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ACT_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed, code " + std::to_string(static_cast<int>(r)));
  return 0;
}

static int make_tmap_2d(CUtensorMap* m, const bf16* ptr, long long inner, long long outer, long long ld, int box_outer) {
  EncodeTiledFn enc = get_encode();
  ACT_CHECK(enc != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  cuuint64_t gd[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t gs[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t bx[2] = {64, static_cast<cuuint32_t>(box_outer)}, es[2] = {1, 1};
  ACT_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && gs[0] % 16 == 0, "TMA 2D alignment");
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<bf16*>(ptr), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  ACT_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2d) failed, code " + std::to_string(static_cast<int>(r)));
  return 0;
}

static int mat_tmap(CUtensorMap* m, const MatDesc& d, int rows, int K, int nz0, int nz1, int box_rows_kmajor) {
  long long dims[4], str[3];
  int box[4] = {64, 64, 1, 1};
  if (d.mn_major) { dims[0] = rows; dims[1] = K; }
  else { dims[0] = K; dims[1] = rows; box[1] = box_rows_kmajor; }
  dims[2] = nz0; dims[3] = nz1;
  str[0] = d.ld;
  str[1] = nz0 > 1 ? d.sz0 : d.ld * std::max(1LL, dims[1]);
  str[2] = nz1 > 1 ? d.sz1 : str[1] * std::max(1, nz0);
  if (str[1] <= 0) str[1] = 8;
  if (str[2] <= 0) str[2] = 8;
  return make_tmap(m, d.ptr, dims, str, box);
}

static int plane_tmap(CUtensorMap* m, const PlaneDesc& pl, int bw, int bh, int bni) {
  long long dims[4] = {pl.C, pl.W, pl.H, pl.N};
  long long str[3] = {pl.sW, pl.sH, pl.sN};
  int box[4] = {64, bw, bh, bni};
  return make_tmap(m, pl.ptr, dims, str, box);
}

// ---- optional per-launch CUDA-event timing of the tensor-core kernels (bench.py roofline leg)
struct TcProf { bool on = false; std::vector<cudaEvent_t> ev; std::vector<double> flops; std::vector<std::string> label; };
static TcProf g_prof;
static double g_next_flops = 0.0;
static std::string g_next_label;
static std::string g_dump_path;
void tc_profile_set_dump(const char* path) { g_dump_path = path ? path : ""; }
void tc_profile_begin() {
  for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
  g_prof.ev.clear(); g_prof.flops.clear(); g_prof.label.clear(); g_prof.on = true;
}
int tc_profile_end(double* total_ms, double* total_flops, int* launches, double* max_ms) {
  g_prof.on = false;
  double ms = 0, fl = 0, mx = 0;
  const int n = static_cast<int>(g_prof.flops.size());
  FILE* f = g_dump_path.empty() ? nullptr : fopen(g_dump_path.c_str(), "w");
  if (f) fprintf(f, "idx,label,ms,gflop,tflops\n");
  for (int i = 0; i < n; ++i) {
    float t = 0.f;
    if (cudaEventSynchronize(g_prof.ev[2 * i + 1]) != cudaSuccess || cudaEventElapsedTime(&t, g_prof.ev[2 * i], g_prof.ev[2 * i + 1]) != cudaSuccess) {
      set_error("tc_profile_end: event query failed"); return -1;
    }
    ms += t; fl += g_prof.flops[i]; if (t > mx) mx = t;
    if (f) fprintf(f, "%d,%s,%.5f,%.4f,%.2f\n", i, g_prof.label[i].c_str(), t, g_prof.flops[i] / 1e9, g_prof.flops[i] / (t * 1e9 + 1e-9));
  }
  if (f) fclose(f);
  for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
  g_prof.ev.clear(); g_prof.flops.clear();
  *total_ms = ms; *total_flops = fl; *launches = n; *max_ms = mx;
  return 0;
}

// Measurement switch (act_tc_set_skip_mask): launches of the selected kernel classes are skipped (their FLOPs are still
// counted), so "serial step time with the class" minus "without it" gives the class's device time from two CUDA-graph
// replays -- no per-launch event pair in the measurement.  Results of a step run this way are garbage by design.
static uint32_t g_skip_mask = 0;
static double g_skipped_flops = 0.0;
enum { CLS_GEMM = 1, CLS_CONV = 2, CLS_CONV3 = 4, CLS_WGRAD = 8, CLS_ATTN = 16 };
void tc_set_skip_mask(uint32_t m) { g_skip_mask = m; g_skipped_flops = 0.0; }
double tc_skipped_flops() { return g_skipped_flops; }
static bool skip_class(uint32_t cls) {
  if (!(g_skip_mask & cls)) return false;
  g_skipped_flops += g_next_flops;
  return true;
}
static int pdl_late() { static int late = -1; if (late < 0) { const char* e = getenv("ACT_PDL_LATE"); late = (e && e[0] == '0') ? 0 : 1; } return late; }
static unsigned long long* g_ts = nullptr;
void tc_debug_timestamps(unsigned long long* buf) { g_ts = buf; }
static int g_sm_reserve = 0;
static int oversub() { static int k = 0; if (!k) { const char* e = getenv("ACT_TC_OVERSUB"); k = e ? std::max(1, atoi(e)) : 1; } return k; }
// SMs left free by the persistent launches that follow (a concurrent latency-bound branch runs there); 0 = use all.
void tc_set_sm_reserve(int n) { g_sm_reserve = n < 0 ? 0 : n; }
static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) n = v;
    else n = 148;
  }
  return std::max(n / 2, n - g_sm_reserve);
}

// Persistent launch: `tiles` = logical tile grid (x fastest); the kernel runs min(#tiles, #SMs) CTAs, one per SM
// (the smem ring is sized to own the SM), each walking tiles blockIdx.x, blockIdx.x + gridDim.x, ...
template <int MODE, int A_MN, int B_MN, int BN>
static int launch_inst(dim3 tiles, int stages, cudaStream_t stream, const CUtensorMap* am, const CUtensorMap& bm,
                       TcParams p) {
  using Cfg = TcCfg<MODE, A_MN, B_MN, BN>;
  auto kern = tc_kernel<MODE, A_MN, B_MN, BN>;
  const int max_smem = 227 * 1024;
  const int extra = 1024 + (2 * Cfg::MAX_STAGES + 5) * 8 + 16;
  // resident weights (MODE_CONV, p.b_res): the ring holds A tiles only and the weight operand sits behind it
  const bool bres = MODE == MODE_CONV && p.b_res != 0;
  const int stage_bytes = bres ? Cfg::A_BYTES : Cfg::STAGE_BYTES;
  const int res_bytes = bres ? p.num_k_blocks * BN * 128 : 0;
  const int fit = std::min<int>(Cfg::MAX_STAGES, (200 * 1024 - res_bytes) / stage_bytes);
  const int stages_in = stages;
  if (stages <= 0) stages = fit;
  stages = std::max(2, std::min(stages, std::min<int>(Cfg::MAX_STAGES, (max_smem - extra - res_bytes) / stage_bytes)));
  ACT_CHECK(stages * stage_bytes + res_bytes + extra <= max_smem, "resident weight operand does not fit in shared memory");
  // Two CTAs never share an SM when either needs more than half of TMEM (they would serialise on the allocation) or walks
  // several tiles (it wants the whole ring).  Single-round launches with <= 256 TMEM columns (most of the batch-8 step:
  // one tile per CTA, latency-bound) may share (ACT_TC_SHARE=1): a short ring per CTA, two CTAs of different branches
  // per SM, one running MMAs while the other waits for its operands.
  static int share = -1;
  if (share < 0) { const char* e = getenv("ACT_TC_SHARE"); share = (e && e[0] == '1') ? 1 : 0; }
  const long long total_tiles = 1LL * tiles.x * tiles.y * tiles.z;
  const bool sharing = share && !bres && Cfg::TMEM_COLS <= 256 && total_tiles <= num_sms() && stages_in <= 0;
  if (sharing) stages = std::min(stages, std::max(2, (104 * 1024 - extra) / stage_bytes));
  const int smem = sharing ? stages * stage_bytes + extra : std::max(stages * stage_bytes + res_bytes + extra, 116 * 1024);
  static bool configured = false;     // per instantiation
  if (!configured) {
    ACT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
    configured = true;
  }
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("ACT_TC_DEBUG"); dbg = e ? atoi(e) : 0; } p.debug = dbg; }
  p.gx = static_cast<int>(tiles.x); p.gy = static_cast<int>(tiles.y); p.gz = static_cast<int>(tiles.z);
  p.ts = g_ts;
  p.pdl_late = pdl_late();
  const long long total = 1LL * tiles.x * tiles.y * tiles.z;
  ACT_CHECK(total > 0 && total < (1LL << 30), "bad tile grid");
  const unsigned ctas = static_cast<unsigned>(std::min<long long>(total, 1LL * oversub() * num_sms()));
  if (skip_class(MODE == MODE_GEMM ? CLS_GEMM : MODE == MODE_CONV ? CLS_CONV : CLS_WGRAD)) return 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_prof.on) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, stream); }
  ACT_CUDA(launch_pdl(kern, dim3(ctas), dim3(sharing ? 192 : Cfg::THREADS), static_cast<size_t>(smem), stream, am[0], am[1], am[2], am[3], bm, p, stages));
  if (g_prof.on) { cudaEventRecord(e1, stream); g_prof.ev.push_back(e0); g_prof.ev.push_back(e1); g_prof.flops.push_back(g_next_flops); g_prof.label.push_back(g_next_label); }
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// Tile width for a persistent launch.  Measured on B200 (tools/bench_kernels.py floor, round 2: the issue loop no longer
// bounds it): one M=128 tcgen05.mma costs 28 / 38 / 74 ns at N = 64 / 128 / 256 (ideal 16 / 33 / 65), and a full-chip
// launch is bounded by L2->SM fill (~85 GB/s per SM when all 148 load), so wide tiles win unless they leave SMs idle.
// cost = rounds x (k-blocks x ns/k-block + tile overhead); ties go to the wider tile.
static int pick_bn(int N, long long row_tiles, int kblocks) {
  int best = 64;
  double best_cost = 1e30;
  const int cands[3] = {64, 128, 256};
  for (int c : cands) {
    if (c > 64 && N < c) continue;
    const long long tiles = row_tiles * ((N + c - 1) / c);
    const long long rounds = (tiles + num_sms() - 1) / num_sms();
    const double active = std::min<double>(1.0, double(tiles) / num_sms());
    const double kb = 16.0 + c / 8.0;                                   // KB of smem fill per k-block
    const double load_ns = kb * std::max(5.0, 12.0 * active);
    const double mma_ns = c == 256 ? 296.0 : c == 128 ? 152.0 : 112.0;
    const double cost = static_cast<double>(rounds) * (kblocks * std::max(load_ns, mma_ns) + 800.0 + 6.0 * c);
    if (cost <= best_cost * 1.0001) { best_cost = std::min(cost, best_cost); best = c; }
  }
  return best;
}

// Split-K plan for a reduction-heavy GEMM (weight gradients: K = tokens): tile width and split minimising
// rounds x (k-blocks per split x ns/k-block + atomic epilogue) on a persistent launch.
void plan_splitk(int M, int N, int K, int* bn_out, int* ks_out) {
  const int mt = (M + 127) / 128, kb = (K + 63) / 64;
  double best = 1e30;
  *bn_out = 64; *ks_out = 1;
  const int cands[3] = {64, 128, 256};
  for (int c : cands) {
    if (c > 64 && N < c) continue;
    const long long base = 1LL * mt * ((N + c - 1) / c);
    const double t_epi = 1500.0 + 12.0 * c;
    for (int k = 1; k <= std::min(64, kb); ++k) {
      const long long tiles = base * k;
      const long long rounds = (tiles + num_sms() - 1) / num_sms();
      const double active = std::min<double>(1.0, double(tiles) / num_sms());
      const double t_kb = std::max((16.0 + c / 8.0) * std::max(5.0, 12.0 * active), c == 256 ? 296.0 : c == 128 ? 152.0 : 112.0);
      const double cost = rounds * (((kb + k - 1) / k) * t_kb + t_epi);
      if (cost < best * 0.999) { best = cost; *bn_out = c; *ks_out = k; }
    }
  }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }
// NHWC outputs: every pixel offset is a multiple of og.ld (and of og.sn), so 32-byte stores need ld % 16 (bf16)
static bool conv_out32(const ConvLaunch& c) {
  return aligned32(c.epi.out_f32) && aligned32(c.epi.out_bf16) && c.og.ld % 16 == 0 && c.og.sn % 16 == 0;
}

static void fill_epi(TcParams& p, const EpiDesc& e) {
  p.flags = e.flags; p.alpha = e.alpha; p.bias = e.bias; p.scale = e.scale; p.res_f32 = e.res_f32;
  p.res_bf16 = e.res_bf16; p.aux_bf16 = e.aux_bf16; p.out_f32 = e.out_f32; p.out_bf16 = e.out_bf16;
  p.ldc = e.ldc; p.ldr = e.ldr; p.osz0 = e.osz0; p.osz1 = e.osz1; p.rsz0 = e.rsz0; p.rsz1 = e.rsz1;
  p.drop = e.drop;                       // (TcParams is memset to 0 by the callers, which is NOT the "off" state)
}

static int check_epi(const EpiDesc& e) {
  ACT_CHECK((e.flags & (EPI_OUT_F32 | EPI_OUT_BF16)) != 0, "epilogue without output");
  if (e.flags & EPI_OUT_F32) ACT_CHECK(e.out_f32 != nullptr, "out_f32 null");
  if (e.flags & EPI_OUT_BF16) ACT_CHECK(e.out_bf16 != nullptr, "out_bf16 null");
  if (e.flags & EPI_BIAS) ACT_CHECK(e.bias != nullptr, "bias null");
  if (e.flags & EPI_SCALE) ACT_CHECK(e.scale != nullptr, "scale null");
  if (e.flags & EPI_RES_F32) ACT_CHECK(e.res_f32 != nullptr, "res_f32 null");
  if (e.flags & EPI_RES_BF16) ACT_CHECK(e.res_bf16 != nullptr, "res_bf16 null");
  if (e.flags & EPI_MASK_POS) ACT_CHECK(e.aux_bf16 != nullptr, "aux null");
  if (e.flags & EPI_ATOMIC) ACT_CHECK((e.flags & EPI_OUT_F32) && !(e.flags & EPI_OUT_BF16), "atomic epilogue is fp32 only");
  return 0;
}

// --------------------------------------------------------------------------------------------- GEMM
int tc_gemm(int M, int N, int K, int nz0, int nz1, const MatDesc& A, const MatDesc& B, const EpiDesc& epi,
            int ksplit, cudaStream_t stream, int bn_hint, int stages_hint) {
  ACT_CHECK(M > 0 && N > 0 && K > 0 && nz0 > 0 && nz1 > 0, "bad GEMM shape");
  if (check_epi(epi)) return -1;
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 1) ACT_CHECK(epi.flags & EPI_ATOMIC, "split-K needs the atomic epilogue");
  if (ksplit > 1) ACT_CHECK(!(epi.flags & (EPI_RELU | EPI_MASK_POS)), "split-K cannot apply a non-linear epilogue");
  if (epi.flags & EPI_DROPOUT)
    ACT_CHECK(ksplit == 1 && nz0 * nz1 == 1 && N % 32 == 0 && epi.drop.step != nullptr && drop_on(epi.drop),
              "dropout epilogue: plain GEMM with N % 32 == 0 and a device step counter only");
  const int nb = nz0 * nz1;
  const int mt = (M + 127) / 128;
  int bn = bn_hint;
  if (bn == 0) bn = pick_bn(N, 1LL * mt * nb * ksplit, std::max(1, ((K + 63) / 64) / ksplit));
  ACT_CHECK(bn == 64 || bn == 128 || bn == 256, "bn must be 64/128/256");
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K; p.num_k_blocks = (K + 63) / 64; p.nz0 = nz0; p.ksplit = ksplit;
  fill_epi(p, epi);
  p.drop_ld = N;
  bool vec = (epi.ldc % 8 == 0) && (epi.osz0 % 8 == 0) && (epi.osz1 % 8 == 0) && (epi.ldr % 8 == 0) &&
             (epi.rsz0 % 8 == 0) && (epi.rsz1 % 8 == 0) && aligned16(epi.out_f32) && aligned16(epi.out_bf16) &&
             aligned16(epi.res_f32) && aligned16(epi.res_bf16) && aligned16(epi.aux_bf16);
  p.vec_ok = vec ? 1 : 0;
  p.vec32_ok = vec && aligned32(epi.out_f32) && aligned32(epi.out_bf16) &&
               (!epi.out_bf16 || (epi.ldc % 16 == 0 && epi.osz0 % 16 == 0 && epi.osz1 % 16 == 0));
  CUtensorMap am[4], bm;
  if (mat_tmap(&am[0], A, M, K, nz0, nz1, 128)) return -1;
  am[1] = am[2] = am[3] = am[0];
  if (mat_tmap(&bm, B, N, K, nz0, nz1, bn)) return -1;
  dim3 grid((N + bn - 1) / bn, mt, nb * ksplit);
  g_next_flops = 2.0 * M * N * K * nb;
  if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "gemm M=%d N=%d K=%d z=%d amn=%d bmn=%d bn=%d ks=%d grid=%ux%ux%u", M, N, K, nb, A.mn_major, B.mn_major, bn, ksplit, grid.x, grid.y, grid.z); g_next_label = buf; }
  const int st = stages_hint;
  const int key = (A.mn_major ? 2 : 0) | (B.mn_major ? 1 : 0);
#define GEMM_CASE(AM, BMJ)                                                                     \
  if (bn == 64) return launch_inst<MODE_GEMM, AM, BMJ, 64>(grid, st, stream, am, bm, p);       \
  if (bn == 128) return launch_inst<MODE_GEMM, AM, BMJ, 128>(grid, st, stream, am, bm, p);     \
  return launch_inst<MODE_GEMM, AM, BMJ, 256>(grid, st, stream, am, bm, p);
  switch (key) {
    case 0: GEMM_CASE(0, 0)
    case 1: GEMM_CASE(0, 1)
    case 2: GEMM_CASE(1, 0)
    default: GEMM_CASE(1, 1)
  }
#undef GEMM_CASE
}

// --------------------------------------------------------------------------------------------- tiles
// Pick the pixel box (BW,BH,BNI), BW*BH*BNI <= 128, maximising useful accumulator rows.
static void pick_tile(int vW, int vH, int nimg, int mult, int max_ni, int* BW, int* BH, int* BNI) {
  double best = -1;
  *BW = 1; *BH = 1; *BNI = 1;
  // boxes may overhang the plane (TMA zero-fills); needed when the pixel count must be a multiple of `mult`
  const int maxw = (mult > 1) ? std::max(vW, mult) : vW, maxh = (mult > 1) ? std::max(vH, mult) : vH;
  for (int bw = 1; bw <= std::min(maxw, 128); ++bw)
    for (int bh = 1; bh * bw <= 128 && bh <= maxh; ++bh)
      for (int bni = 1; bni * bh * bw <= 128 && bni <= std::min(nimg, max_ni); ++bni) {
        if ((bw * bh * bni) % mult) continue;
        if (bni > 1 && (bh != vH || bw != vW)) continue;      // span images only with whole planes
        const long long tiles = 1LL * ((vW + bw - 1) / bw) * ((vH + bh - 1) / bh) * ((nimg + bni - 1) / bni);
        double util = double(vW) * vH * nimg / (double(tiles) * 128.0);
        util += 1e-4 * bw;                                      // prefer long contiguous rows
        if (util > best) { best = util; *BW = bw; *BH = bh; *BNI = bni; }
      }
}

// --------------------------------------------------------------------------------------------- CONV
int tc_conv(const ConvLaunch& c, cudaStream_t stream, int bn_hint, int stages_hint) {
  if (tc_conv3_applicable(c)) return tc_conv3(c, stream, bn_hint);
  ACT_CHECK(c.Cin % 64 == 0 && c.Cin >= 64, "Cin must be a multiple of 64");
  ACT_CHECK(c.ntaps >= 1 && c.ntaps <= 16 && c.nplanes >= 1 && c.nplanes <= 4, "bad tap table");
  if (check_epi(c.epi)) return -1;
  TcParams p;
  memset(&p, 0, sizeof(p));
  int BW, BH, BNI;
  pick_tile(c.vW, c.vH, c.NIMG, 1, 256, &BW, &BH, &BNI);
  p.BW = BW; p.BH = BH; p.BNI = BNI;
  p.tiles_w = (c.vW + BW - 1) / BW; p.tiles_h = (c.vH + BH - 1) / BH;
  const int tiles_n = (c.NIMG + BNI - 1) / BNI;
  p.NIMG = c.NIMG; p.vH = c.vH; p.vW = c.vW; p.og = c.og; p.rg = c.rg;
  p.ntaps = c.ntaps; p.cin_blocks = c.Cin / 64; p.num_k_blocks = c.ntaps * p.cin_blocks;
  p.N = c.Cout; p.nz0 = 1; p.ksplit = 1;
  for (int i = 0; i < c.ntaps; ++i) p.taps[i] = c.taps[i];
  fill_epi(p, c.epi);
  p.vec_ok = (c.og.ld % 8 == 0) && (c.rg.ld % 8 == 0 || !(c.epi.flags & (EPI_RES_F32 | EPI_RES_BF16))) &&
             aligned16(c.epi.out_f32) && aligned16(c.epi.out_bf16) && aligned16(c.epi.res_f32) &&
             aligned16(c.epi.res_bf16) && aligned16(c.epi.aux_bf16);
  p.vec32_ok = p.vec_ok && conv_out32(c);
  const long long mtiles = 1LL * p.tiles_w * p.tiles_h * tiles_n;
  int bn = bn_hint;
  if (bn == 0) bn = pick_bn(c.Cout, mtiles, c.ntaps * (c.Cin / 64));
  CUtensorMap am[4], bm;
  for (int i = 0; i < 4; ++i)
    if (plane_tmap(&am[i], c.planes[std::min(i, c.nplanes - 1)], BW, BH, BNI)) return -1;
  if (make_tmap_2d(&bm, c.wgt, c.wgt_ld, c.Cout, c.wgt_ld, bn)) return -1;
  dim3 grid((c.Cout + bn - 1) / bn, static_cast<unsigned>(mtiles), 1);
  {
    // Weights resident in shared memory when one N tile covers Cout, they fit beside >= 4 A stages, and every CTA walks
    // enough tiles to amortise the up-front load (ACT_CONV_BRES=0 disables).
    static int on = -1;
    if (on < 0) { const char* e = getenv("ACT_CONV_BRES"); on = (e && e[0] == '0') ? 0 : 1; }
    const long long res = 1LL * p.num_k_blocks * bn * 128;
    p.b_res = (on && grid.x == 1 && res <= 128 * 1024 && mtiles >= 4LL * num_sms()) ? 1 : 0;
  }
  g_next_flops = c.algo_flops > 0 ? c.algo_flops : 2.0 * c.vW * c.vH * c.NIMG * c.Cout * c.Cin * c.ntaps;
  if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "conv %dx%dx%d Cin=%d Cout=%d taps=%d box=%dx%dx%d bn=%d grid=%ux%u%s", c.NIMG, c.vH, c.vW, c.Cin, c.Cout, c.ntaps, BNI, BH, BW, bn, grid.x, grid.y, p.b_res ? " wres" : ""); g_next_label = buf; }
  const int st = stages_hint;
  if (bn == 64) return launch_inst<MODE_CONV, 0, 0, 64>(grid, st, stream, am, bm, p);
  if (bn == 128) return launch_inst<MODE_CONV, 0, 0, 128>(grid, st, stream, am, bm, p);
  return launch_inst<MODE_CONV, 0, 0, 256>(grid, st, stream, am, bm, p);
}

// --------------------------------------------------------------------------------------------- CONV3 (halo-resident)
template <int BN>
static int launch_conv3(long long total, cudaStream_t stream, const CUtensorMap& am, const CUtensorMap& bm, const TcParams& p) {
  using Cfg = Conv3Cfg<BN>;
  auto kern = conv3_kernel<BN>;
  static bool configured = false;
  if (!configured) {
    ACT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  const unsigned ctas = static_cast<unsigned>(std::min<long long>(total, 1LL * oversub() * num_sms()));
  if (skip_class(CLS_CONV3)) return 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_prof.on) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, stream); }
  ACT_CUDA(launch_pdl(kern, dim3(ctas), dim3(320), static_cast<size_t>(Cfg::SMEM), stream, am, bm, p));
  if (g_prof.on) { cudaEventRecord(e1, stream); g_prof.ev.push_back(e0); g_prof.ev.push_back(e1); g_prof.flops.push_back(g_next_flops); g_prof.label.push_back(g_next_label); }
  ACT_CUDA(cudaGetLastError());
  return 0;
}

bool tc_conv3_applicable(const ConvLaunch& c) {
  static int on = -1;
  if (on < 0) { const char* e = getenv("ACT_CONV3"); on = (e && e[0] == '0') ? 0 : 1; }
  if (!on || c.nplanes != 1 || c.ntaps < 1 || c.ntaps > 9 || c.Cin % 64 || c.Cout % 64) return false;
  for (int i = 0; i < c.ntaps; ++i)
    if (c.taps[i].map != 0 || c.taps[i].dh < -1 || c.taps[i].dh > 1 || c.taps[i].dw < -1 || c.taps[i].dw > 1) return false;
  if (c.ntaps < 4) return false;                       // 1x1 / sparse tap sets gain nothing from the resident patch
  const PlaneDesc& pl = c.planes[0];
  return pl.W == c.vW && pl.H == c.vH && pl.N == c.NIMG;
}

int tc_conv3(const ConvLaunch& c, cudaStream_t stream, int bn_hint) {
  ACT_CHECK(tc_conv3_applicable(c), "tc_conv3: not a dense 3x3 / stride-1 tap set");
  if (check_epi(c.epi)) return -1;
  TcParams p;
  memset(&p, 0, sizeof(p));
  // tile geometry: BH x BW outputs per 128-row accumulator, patch pitch PW = BW + 2 (<= 46 so every tap window fits)
  double best = -1;
  int BW = 1, BH = 1;
  for (int bw = 2; bw <= std::min(c.vW, 44); ++bw) {
    const int pw = bw + 2;
    int bh = std::min(c.vH, 128 / pw);
    while (bh > 1 && (bh + 2) * pw > 222) --bh;
    if (bh < 1 || (bh + 2) * pw > 222) continue;
    const long long tiles = 1LL * ((c.vW + bw - 1) / bw) * ((c.vH + bh - 1) / bh);
    const double util = double(c.vW) * c.vH / (double(tiles) * 128.0) + 1e-4 * bw;
    if (util > best) { best = util; BW = bw; BH = bh; }
  }
  p.BW = BW; p.BH = BH; p.PW = BW + 2; p.BNI = 1;
  p.tiles_w = (c.vW + BW - 1) / BW; p.tiles_h = (c.vH + BH - 1) / BH;
  p.NIMG = c.NIMG; p.vH = c.vH; p.vW = c.vW; p.og = c.og; p.rg = c.rg;
  p.ntaps = c.ntaps; p.cin_blocks = c.Cin / 64; p.N = c.Cout; p.nz0 = 1; p.ksplit = 1;
  p.skip_mask = c.skip_mask;
  ACT_CHECK(c.skip_mask == 0 || c.ntaps * (c.Cin / 64) <= 32, "tc_conv3: skip mask covers at most 32 weight blocks");
  for (int i = 0; i < c.ntaps; ++i) p.taps[i] = c.taps[i];
  fill_epi(p, c.epi);
  p.vec_ok = (c.og.ld % 8 == 0) && (c.rg.ld % 8 == 0 || !(c.epi.flags & (EPI_RES_F32 | EPI_RES_BF16))) &&
             aligned16(c.epi.out_f32) && aligned16(c.epi.out_bf16) && aligned16(c.epi.res_f32) &&
             aligned16(c.epi.res_bf16) && aligned16(c.epi.aux_bf16);
  p.vec32_ok = p.vec_ok && conv_out32(c);
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("ACT_TC_DEBUG"); dbg = e ? atoi(e) : 0; } p.debug = dbg; }
  p.pdl_late = pdl_late();
  const long long mtiles = 1LL * p.tiles_w * p.tiles_h * c.NIMG;
  int bn = bn_hint;
  if (bn == 0) {
    double bc = 1e30;
    const int cands[3] = {64, 128, 256};
    for (int cc : cands) {
      if (c.Cout % cc) continue;
      const long long tiles = mtiles * (c.Cout / cc);
      const long long rounds = (tiles + num_sms() - 1) / num_sms();
      const double cost = rounds * (p.cin_blocks * c.ntaps * 4 * (cc == 256 ? 74.0 : cc == 128 ? 38.0 : 28.0) + 1000.0 + 6.0 * cc);
      if (cost <= bc * 1.0001) { bc = std::min(bc, cost); bn = cc; }
    }
  }
  ACT_CHECK(bn == 64 || bn == 128 || bn == 256, "bn must be 64/128/256");
  ACT_CHECK(c.Cout % bn == 0, "tc_conv3: Cout must be a multiple of the tile width");
  p.gx = c.Cout / bn; p.gy = static_cast<int>(mtiles); p.gz = 1;
  CUtensorMap am, bm;
  if (plane_tmap(&am, c.planes[0], p.PW, BH + 2, 1)) return -1;
  if (make_tmap_2d(&bm, c.wgt, c.wgt_ld, c.Cout, c.wgt_ld, bn)) return -1;
  g_next_flops = c.algo_flops > 0 ? c.algo_flops : 2.0 * c.vW * c.vH * c.NIMG * c.Cout * c.Cin * c.ntaps;
  if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "conv3 %dx%dx%d Cin=%d Cout=%d taps=%d tile=%dx%d bn=%d grid=%dx%d", c.NIMG, c.vH, c.vW, c.Cin, c.Cout, c.ntaps, BH, BW, bn, p.gx, p.gy); g_next_label = buf; }
  const long long total = 1LL * p.gx * p.gy;
  if (bn == 64) return launch_conv3<64>(total, stream, am, bm, p);
  if (bn == 128) return launch_conv3<128>(total, stream, am, bm, p);
  return launch_conv3<256>(total, stream, am, bm, p);
}

// --------------------------------------------------------------------------------------------- WGRAD
int tc_wgrad(const WgradLaunch& w, cudaStream_t stream) {
  ACT_CHECK(w.Cin % 64 == 0 && w.Cout % 64 == 0, "channels must be multiples of 64");
  ACT_CHECK(w.Cin == 64 || w.Cin % 128 == 0, "Cin must be 64 or a multiple of 128");
  ACT_CHECK(w.dw != nullptr && w.ntaps >= 1 && w.ntaps <= 16, "bad wgrad args");
  TcParams p;
  memset(&p, 0, sizeof(p));
  int BW, BH, BNI;
  pick_tile(w.dy.W, w.dy.H, 1, 16, 1, &BW, &BH, &BNI);
  ACT_CHECK((BW * BH) % 16 == 0, "no pixel tile with a multiple of 16 pixels");
  p.BW = BW; p.BH = BH; p.BNI = 1; p.pix_per_box = BW * BH;
  p.tiles_w = (w.dy.W + BW - 1) / BW; p.tiles_h = (w.dy.H + BH - 1) / BH;
  p.num_k_blocks = p.tiles_w * p.tiles_h * w.dy.N;
  p.ntaps = w.ntaps; p.KK = w.KK; p.Cin = w.Cin; p.Cout = w.Cout; p.N = w.Cout;
  p.taps_per_cta = (w.Cin == 64) ? 2 : 1;
  for (int i = 0; i < w.ntaps; ++i) p.taps[i] = w.taps[i];
  p.nz0 = 1; p.alpha = w.alpha; p.out_f32 = w.dw; p.flags = EPI_OUT_F32 | EPI_ATOMIC;
  const int bn = (w.Cout >= 128) ? 128 : 64;
  const int ygroups = (p.taps_per_cta == 2) ? (w.ntaps + 1) / 2 : w.ntaps * (w.Cin / 128);
  const int xt = (w.Cout + bn - 1) / bn;
  int ks = w.ksplit;
  if (ks <= 0) {
    // persistent launch: rounds x (k-blocks per split x MMA time + epilogue); pick the split with the lowest cost
    const long long base = 1LL * ygroups * xt;
    const double t_kb = (p.pix_per_box / 16) * (bn == 128 ? 38.0 : 28.0), t_epi = 2500.0;
    double best = 1e30;
    ks = 1;
    for (int c = 1; c <= std::min(96, p.num_k_blocks); ++c) {
      const long long rounds = (base * c + num_sms() - 1) / num_sms();
      const double cost = rounds * (((p.num_k_blocks + c - 1) / c) * t_kb + t_epi);
      if (cost < best * 0.999) { best = cost; ks = c; }
    }
  }
  ks = std::min(ks, p.num_k_blocks);
  p.ksplit = ks;
  CUtensorMap am[4], bm;
  for (int i = 0; i < 4; ++i)
    if (plane_tmap(&am[i], w.planes[std::min(i, w.nplanes - 1)], BW, BH, 1)) return -1;
  if (plane_tmap(&bm, w.dy, BW, BH, 1)) return -1;
  dim3 grid(xt, ygroups, ks);
  g_next_flops = w.algo_flops > 0 ? w.algo_flops : 2.0 * w.dy.W * w.dy.H * w.dy.N * w.Cout * w.Cin * w.ntaps;
  if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "wgrad %dx%dx%d Cin=%d Cout=%d taps=%d box=%dx%d bn=%d ks=%d grid=%ux%ux%u", w.dy.N, w.dy.H, w.dy.W, w.Cin, w.Cout, w.ntaps, BH, BW, bn, ks, grid.x, grid.y, grid.z); g_next_label = buf; }
  if (bn == 64) return launch_inst<MODE_WGRAD, 1, 1, 64>(grid, 0, stream, am, bm, p);
  return launch_inst<MODE_WGRAD, 1, 1, 128>(grid, 0, stream, am, bm, p);
}

}  // namespace act

namespace act {
// --------------------------------------------------------------------------------------------- fused attention scores
template <int W, int BWD>
static int launch_attn(const CUtensorMap& am, const CUtensorMap& bm, const AttnParams& p, int tiles, cudaStream_t stream) {
  using Cfg = AttnCfg<W>;
  auto kern = attn_scores_kernel<W, BWD>;
  static bool configured = false;
  if (!configured) {
    ACT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  const unsigned ctas = static_cast<unsigned>(std::min(tiles, num_sms()));
  if (skip_class(CLS_ATTN)) return 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_prof.on) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, stream); }
  ACT_CUDA(launch_pdl(kern, dim3(ctas), dim3(256), static_cast<size_t>(Cfg::SMEM), stream, am, bm, p));
  if (g_prof.on) { cudaEventRecord(e1, stream); g_prof.ev.push_back(e0); g_prof.ev.push_back(e1); g_prof.flops.push_back(g_next_flops); g_prof.label.push_back(g_next_label); }
  ACT_CUDA(cudaGetLastError());
  return 0;
}

bool tc_attn_scores_applicable(int Lk, int Lkp, bool need_fused) {
  static int on = -1;
  if (on < 0) { const char* e = getenv("ACT_FUSED_ATTN"); on = !(e && e[0] == '0'); }
  if (Lk > 320 && !need_fused) {                             // (read per call: tests switch it inside one process)
    const char* e = getenv("ACT_ATTN_CHUNKED");
    if (e && e[0] == '0') return false;                      // A/B switch: tc_gemm + softmax for long rows
  }
  return on && Lk >= 1 && Lk <= AttnCCfg::NCH_MAX * AttnCCfg::CH && Lkp % 8 == 0 && Lkp >= Lk;
}

template <int BWD>
static int launch_attn_chunked(const CUtensorMap& am, const CUtensorMap& bm, const AttnParams& p, int items, cudaStream_t stream) {
  auto kern = attn_scores_chunked_kernel<BWD>;
  static bool configured = false;
  if (!configured) {
    ACT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnCCfg::SMEM));
    configured = true;
  }
  const unsigned ctas = static_cast<unsigned>(std::min(items, num_sms()));
  if (skip_class(CLS_ATTN)) return 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (g_prof.on) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, stream); }
  ACT_CUDA(launch_pdl(kern, dim3(ctas), dim3(128 * AttnCCfg::PARTS), static_cast<size_t>(AttnCCfg::SMEM), stream, am, bm, p));
  if (g_prof.on) { cudaEventRecord(e1, stream); g_prof.ev.push_back(e0); g_prof.ev.push_back(e1); g_prof.flops.push_back(g_next_flops); g_prof.label.push_back(g_next_label); }
  ACT_CUDA(cudaGetLastError());
  return 0;
}

// FWD: out = softmax(scale * A B^T + mask) with A = Q [Lq x 64], B = K [Lk x 64] per (head, batch)
// BWD: out = scale * P o (dP - rowsum(dP o P)), dP = A B^T with A = dO, B = V
int tc_attn_scores(int bwd, const MatDesc& A, const MatDesc& B, int Lq, int Lk, int Lkp, int heads, int batch, float scale,
                   const uint8_t* mask, const bf16* P, bf16* out, cudaStream_t stream, const DropDesc* drop, bf16* out2,
                   const MatDesc* O) {
  ACT_CHECK(tc_attn_scores_applicable(Lk, Lkp), "fused attention scores need Lk <= 1024 and a padded stride % 8 == 0");
  ACT_CHECK(!(drop && drop_on(*drop)) || bwd || out2, "fused attention scores: dropout needs the second (dropped) output");
  ACT_CHECK(!(drop && drop_on(*drop)) || drop->step, "fused attention scores: dropout needs the device step counter");
  ACT_CHECK(!A.mn_major && !B.mn_major && out && (!bwd || P), "fused attention scores: bad operands");
  ACT_CHECK((reinterpret_cast<uintptr_t>(out) & 31) == 0 && (!P || (reinterpret_cast<uintptr_t>(P) & 15) == 0), "fused attention scores: unaligned P / dS");
  const int W = Lk <= 128 ? 128 : 320;
  AttnParams p;
  memset(&p, 0, sizeof(p));
  p.Lq = Lq; p.Lk = Lk; p.Lkp = Lkp; p.heads = heads; p.batch = batch; p.qtiles = (Lq + 127) / 128; p.scale = scale;
  p.O = nullptr; p.dO = nullptr; p.o_ld = p.o_sz0 = p.o_sz1 = p.do_ld = p.do_sz0 = p.do_sz1 = 0;
  if (bwd && O && O->ptr && !A.mn_major && (A.ld % 8) == 0 && (O->ld % 8) == 0 && (A.sz0 % 8) == 0 && (O->sz0 % 8) == 0 &&
      (reinterpret_cast<uintptr_t>(A.ptr) & 15) == 0 && (reinterpret_cast<uintptr_t>(O->ptr) & 15) == 0) {   // 16-byte row pieces
    p.O = O->ptr; p.o_ld = O->ld; p.o_sz0 = O->sz0; p.o_sz1 = O->sz1;
    p.dO = A.ptr; p.do_ld = A.ld; p.do_sz0 = A.sz0; p.do_sz1 = A.sz1;
  }
  p.mask = mask; p.P = P; p.out = out; p.vec32 = (Lkp % 16 == 0) ? 1 : 0;
  p.drop = drop ? *drop : DropDesc();            // (the memset above is not the "off" state: thresh 0 would drop everything)
  p.out2 = out2;
  ACT_CHECK(!out2 || (reinterpret_cast<uintptr_t>(out2) & 31) == 0, "fused attention scores: unaligned dropped-P output");
  CUtensorMap am, bm;
  if (mat_tmap(&am, A, Lq, 64, heads, batch, 128)) return -1;
  if (Lk > 320) {
    // chunked two-pass kernel: items = (batch, head, group of query tiles), enough groups to fill the chip
    p.nch = (Lk + AttnCCfg::CH - 1) / AttnCCfg::CH;
    const int pairs = heads * batch;
    p.groups = std::max(1, std::min(p.qtiles, (num_sms() + pairs - 1) / pairs));
    p.tiles_per_group = (p.qtiles + p.groups - 1) / p.groups;
    p.groups = (p.qtiles + p.tiles_per_group - 1) / p.tiles_per_group;
    if (mat_tmap(&bm, B, Lk, 64, heads, batch, AttnCCfg::CH)) return -1;
    g_next_flops = 2.0 * Lq * Lk * 64 * heads * batch;          // algorithmic (the second pass recomputes the scores)
    if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "attn_%s Lq=%d Lk=%d z=%d chunked nch=%d groups=%d", bwd ? "dS" : "P", Lq, Lk, pairs, p.nch, p.groups); g_next_label = buf; }
    return bwd ? launch_attn_chunked<1>(am, bm, p, pairs * p.groups, stream) : launch_attn_chunked<0>(am, bm, p, pairs * p.groups, stream);
  }
  if (mat_tmap(&bm, B, Lk, 64, heads, batch, W == 128 ? 128 : 160)) return -1;
  const int tiles = p.qtiles * heads * batch;
  g_next_flops = 2.0 * Lq * Lk * 64 * heads * batch;
  if (g_prof.on) { char buf[160]; snprintf(buf, sizeof buf, "attn_%s Lq=%d Lk=%d z=%d W=%d", bwd ? "dS" : "P", Lq, Lk, heads * batch, W); g_next_label = buf; }
  if (W == 128) return bwd ? launch_attn<128, 1>(am, bm, p, tiles, stream) : launch_attn<128, 0>(am, bm, p, tiles, stream);
  return bwd ? launch_attn<320, 1>(am, bm, p, tiles, stream) : launch_attn<320, 0>(am, bm, p, tiles, stream);
}
}  // namespace act
