// Host-side (C++) interface of the tcgen05 contraction kernels.  Internal to the library; the public
// C ABI lives in include/act_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string>
#include "tc_kernel.cuh"
#include "dropout.cuh"

namespace act {

typedef __nv_bfloat16 bf16;

void set_error(const std::string& msg);
const char* last_error();

#define ACT_CHECK(cond, msg)                                                         \
  do {                                                                               \
    if (!(cond)) { ::act::set_error(std::string(__FILE__) + ":" + std::to_string(__LINE__) + " " + (msg)); return -1; } \
  } while (0)
#define ACT_CUDA(expr)                                                               \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) { ::act::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); return -2; } \
  } while (0)

// Launch with the programmatic-stream-serialization attribute (PDL); ACT_PDL=0 in the environment turns it off.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Logical matrix operand [rows x K] (batched).  K-major: (row,k) at ptr + row*ld + k.
// MN-major: (row,k) at ptr + k*ld + row.  Batch z = z0 + nz0*z1 adds z0*sz0 + z1*sz1 elements.
struct MatDesc {
  const bf16* ptr = nullptr;
  int mn_major = 0;
  long long ld = 0, sz0 = 0, sz1 = 0;
};

struct EpiDesc {
  uint32_t flags = 0;
  float alpha = 1.f;
  const float* bias = nullptr;
  const float* scale = nullptr;
  const float* res_f32 = nullptr;
  const bf16* res_bf16 = nullptr;
  const bf16* aux_bf16 = nullptr;
  float* out_f32 = nullptr;
  bf16* out_bf16 = nullptr;
  long long ldc = 0, ldr = 0, osz0 = 0, osz1 = 0, rsz0 = 0, rsz1 = 0;
  DropDesc drop;                        // EPI_DROPOUT (plain GEMMs only): keep mask indexed row * N + column
};

// D[z][M,N] = epi(A[z] * B[z]^T).  bn_hint / stages_hint = 0 -> heuristic.
int tc_gemm(int M, int N, int K, int nz0, int nz1, const MatDesc& A, const MatDesc& B, const EpiDesc& epi,
            int ksplit, cudaStream_t stream, int bn_hint = 0, int stages_hint = 0);

void plan_splitk(int M, int N, int K, int* bn, int* ksplit);   // tile width + split for reduction-heavy GEMMs

// One NHWC plane (possibly a stride-2 parity sub-lattice) addressed by TMA: dims {C, W, H, N}.
struct PlaneDesc {
  const bf16* ptr = nullptr;
  int C = 0, W = 0, H = 0, N = 0;
  long long sW = 0, sH = 0, sN = 0;     // element strides (channel stride is 1)
};

struct ConvLaunch {
  PlaneDesc planes[4];
  int nplanes = 1;
  TapDesc taps[16];
  int ntaps = 0;
  int Cin = 0, Cout = 0;
  const bf16* wgt = nullptr;            // [Cout][nslots*Cin] K-major, slot-major then channel
  long long wgt_ld = 0;
  int vW = 0, vH = 0, NIMG = 0;         // tile-space extent (one accumulator row per (n,h,w))
  PixGeom og{}, rg{};
  EpiDesc epi;                          // ldc/ldr/osz*/rsz* ignored (geometry comes from og / rg)
  double algo_flops = 0.0;              // > 0: algorithmic FLOPs of this launch for the profile (operands with zero padding in K)
  uint32_t skip_mask = 0;               // dense 3x3 tap sets on conv3_kernel: bit (tap index * Cin/64 + channel block) = zero block
};
int tc_conv(const ConvLaunch& c, cudaStream_t stream, int bn_hint = 0, int stages_hint = 0);
// dense 3x3 / stride-1 tap sets: halo patch resident in shared memory (conv3_kernel.cuh); ACT_CONV3=0 disables it
bool tc_conv3_applicable(const ConvLaunch& c);
void tc_set_sm_reserve(int n);
// Attention scores fused with softmax (bwd = 0) / softmax backward (bwd = 1) for head_dim 64 and Lk <= 1024 keys
// (one TMEM-resident pass up to 320 keys, two chunked passes beyond).  drop (optional, train mode): dropout on the
// attention weights -- forward also writes out2 = dropout(P), backward masks dP (csrc/dropout.cuh).
// need_fused: the caller cannot use the unfused pair (attention dropout lives in the fused kernels only).  Rows of more than
// 320 keys take the chunked kernel (210 / 258 us fwd / bwd at 16 x 8 x 902 x 902 against 210 / 297 us for tc_gemm + softmax,
// profiles/r02_attn_902.md); ACT_ATTN_CHUNKED=0 routes them to the unfused pair unless need_fused.
bool tc_attn_scores_applicable(int Lk, int Lkp, bool need_fused = true);
int tc_attn_scores(int bwd, const MatDesc& A, const MatDesc& B, int Lq, int Lk, int Lkp, int heads, int batch, float scale,
                   const uint8_t* mask, const bf16* P, bf16* out, cudaStream_t stream, const DropDesc* drop = nullptr,
                   bf16* out2 = nullptr, const MatDesc* O = nullptr);   // O (bwd, optional): the forward's attention output
   // persistent launches that follow leave n SMs free for a concurrent branch
int tc_conv3(const ConvLaunch& c, cudaStream_t stream, int bn_hint = 0);

struct WgradLaunch {
  PlaneDesc planes[4];                  // activation planes (tap.map selects)
  int nplanes = 1;
  TapDesc taps[16];
  int ntaps = 0;
  PlaneDesc dy;                         // gradient wrt the conv output, dims {Cout, vW, vH, N}
  int Cin = 0, Cout = 0, KK = 0;        // staged output T fp32 [KK][Cout][Cin] (ci contiguous), accumulated with
  float* dw = nullptr;                  // vector atomics (caller zeroes it and folds it into dW[Cout][Cin][KK])
  float alpha = 1.f;
  const float* scale = nullptr;         // unused by the kernel (applied by the fold)
  int ksplit = 0;                       // 0 -> heuristic
  double algo_flops = 0.0;              // > 0: algorithmic FLOPs of this launch for the profile
};
int tc_wgrad(const WgradLaunch& w, cudaStream_t stream);

// CUDA-event timing of every tensor-core launch between begin and end (algorithmic FLOPs summed per launch)
void tc_debug_timestamps(unsigned long long* buf);   // diagnostics: CTA 0 of every tc_kernel launch stamps %globaltimer into buf
void tc_set_skip_mask(uint32_t mask);         // measurement: skip launches of kernel classes (1 gemm, 2 conv, 4 conv3, 8 wgrad, 16 attention)
double tc_skipped_flops();                    // algorithmic FLOPs of the launches skipped since the mask was set
void tc_profile_begin();
void tc_profile_set_dump(const char* path);   // per-launch CSV written by tc_profile_end (empty: off)
int tc_profile_end(double* total_ms, double* total_flops, int* launches, double* max_ms);

}  // namespace act
