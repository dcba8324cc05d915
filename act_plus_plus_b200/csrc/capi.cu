// extern "C" surface of libact_b200.so (declared in include/act_b200.h).
#include "../../include/act_b200.h"
#include "tc_api.h"
#include "conv_ops.h"
#include "kernels.h"

using namespace act;

extern "C" const char* act_last_error(void) { return last_error(); }
extern "C" int act_abi_version(void) { return 1; }

static EpiDesc to_epi(const act_epilogue* e) {
  EpiDesc d;
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
