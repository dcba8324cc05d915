// extern "C" surface of libact_b200.so (declared in include/act_b200.h).
#include "../../include/act_b200.h"
#include "tc_api.h"

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
