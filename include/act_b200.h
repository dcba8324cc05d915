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

#ifdef __cplusplus
}
#endif
#endif /* ACT_B200_H_ */
