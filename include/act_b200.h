/* act_b200 -- C ABI of the B200-native ACT (DETRVAE) hot path.
 *
 * Plain C: raw device pointers, sizes, an explicit cudaStream_t passed as void*.  Every function returns 0 on
 * success and a negative code on failure; act_last_error() gives the message (no exceptions cross the ABI).
 * The reference (jie0530/act-plus-plus) is pure Python/PyTorch and has no FFI; each entry point below names the
 * reference code (file:line under /root/reference) whose work it replaces.  INTEGRATION.md shows the ctypes
 * binding a maintainer of the reference would add.
 *
 * Layout conventions: bf16 tensors are raw uint16 storage; activations are NHWC; token tensors are [B, L, d].
 */
#ifndef ACT_B200_H_
#define ACT_B200_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* act_last_error(void);
int act_abi_version(void);

/* ---- dense contractions (tcgen05 / TMA).  Replace torch nn.Linear / bmm calls of transformer.py:194-259 and
 *      detr_vae.py:57-63 (F.linear -> cuBLAS in the reference). ---------------------------------------------- */
typedef struct {
  uint32_t flags;            /* ACT_EPI_* */
  float alpha;
  const float* bias;         /* [N] */
  const float* scale;        /* [N] */
  const float* res_f32;
  const void* res_bf16;
  const void* aux_bf16;
  float* out_f32;
  void* out_bf16;
  int64_t ldc, ldr, osz0, osz1, rsz0, rsz1;
} act_epilogue;

enum {
  ACT_EPI_BIAS = 1, ACT_EPI_SCALE = 2, ACT_EPI_RES_F32 = 4, ACT_EPI_RES_BF16 = 8, ACT_EPI_RELU = 16,
  ACT_EPI_MASK_POS = 32, ACT_EPI_OUT_F32 = 64, ACT_EPI_OUT_BF16 = 128, ACT_EPI_ATOMIC = 256
};

/* D[z][M,N] = epilogue(alpha * A[z][M,K] * B[z][N,K]^T), bf16 operands, fp32 accumulation in TMEM.
 * a_mn / b_mn = 0: operand stored K-major ((row,k) at ptr + row*ld + k); 1: MN-major ((row,k) at ptr + k*ld + row).
 * Batch z = z0 + nz0*z1 adds z0*sz0 + z1*sz1 elements to each operand. */
int act_tc_gemm(int M, int N, int K, int nz0, int nz1,
                const void* A, int a_mn, int64_t lda, int64_t a_sz0, int64_t a_sz1,
                const void* B, int b_mn, int64_t ldb, int64_t b_sz0, int64_t b_sz1,
                const act_epilogue* epi, int ksplit, int bn_hint, int stages_hint, void* stream);


/* ---- convolutions over NHWC bf16 (implicit GEMM on tcgen05, TMA tap boxes).  Replace torchvision resnet18's
 *      nn.Conv2d + FrozenBatchNorm2d (backbone.py:47-57,121-124) and their autograd backward (cuDNN in the
 *      reference).  Kernels are 1x1 or 3x3, stride 1 or 2.
 *      wpk  : bf16 [Cout][(kh*KW+kw)*Cin + ci]           (forward operand)
 *      wpkt : bf16 [Cin][(kh*KW+kw)*Cout + co] * scale   (data-gradient operand)                         ---- */
int act_cast_bf16(const float* src, void* dst, int64_t n, void* stream);
int act_pack_conv_weights(const float* w, const float* scale, void* wpk, void* wpkt, int Cout, int Cin, int KK,
                          void* stream);
/* y[N,Ho,Wo,Cout] = epilogue(conv(x[N,H,W,Cin], w)); epilogue geometry (ldc etc.) is implied by the shape. */
int act_conv_fwd(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x, const void* wpk,
                 const act_epilogue* epi, void* stream);
/* dx[N,H,W,Cin] = epilogue(conv_transpose(dy[N,Ho,Wo,Cout], w)); for stride 2 `res_compact` ([N,H/2,W/2,Cin] bf16,
 * may be NULL) is added on the even/even pixels and epi's own residual is ignored. */
int act_conv_dgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* dy,
                   const void* wpkt, const act_epilogue* epi, const void* res_compact, void* stream);
/* dw[Cout][Cin][KH][KW] (fp32) += alpha * scale[co] * sum_pixels dy * x  (atomic accumulation). */
int act_conv_wgrad(int N, int H, int W, int Cin, int Cout, int KH, int stride, int pad, const void* x, const void* dy,
                   float* dw, float alpha, const float* scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ACT_B200_H_ */
