// One warp-specialised tcgen05 kernel template shared by every dense contraction on the ACT path:
//   MODE_GEMM  : batched D[z][M,N] = A[z][M,K] * B[z][N,K]^T, each operand K-major or MN-major
//   MODE_CONV  : implicit-GEMM convolution / transposed convolution over NHWC bf16 (tap table, TMA boxes)
//   MODE_WGRAD : convolution weight gradient, reduction over pixels, both operands MN-major, split-K atomics
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> global).
// Operands land in 128B-swizzled shared memory via TMA; accumulators live in TMEM (fp32).
#pragma once
#include "ptx.cuh"

namespace act {

enum { MODE_GEMM = 0, MODE_CONV = 1, MODE_WGRAD = 2 };

enum EpiFlags : uint32_t {
  EPI_BIAS = 1u << 0,        // + bias[col]                       (fp32)
  EPI_SCALE = 1u << 1,       // * scale[col]   (applied before bias; FrozenBN fold)
  EPI_RES_F32 = 1u << 2,     // + res_f32[addr]
  EPI_RES_BF16 = 1u << 3,    // + res_bf16[addr]
  EPI_RELU = 1u << 4,        // max(.,0)
  EPI_MASK_POS = 1u << 5,    // zero where aux_bf16[addr] <= 0   (ReLU backward against the saved output)
  EPI_OUT_F32 = 1u << 6,
  EPI_OUT_BF16 = 1u << 7,
  EPI_ATOMIC = 1u << 8,      // fp32 atomicAdd instead of store (split-K)
  EPI_ROWSCALE = 1u << 9,    // * rowscale[z*rs_ld + row]  (unused hooks keep the struct stable)
};

struct TapDesc {     // one filter tap of an implicit-GEMM convolution
  int16_t map;       // which A tensor map (stride-2 parity plane)
  int16_t dh, dw;    // coordinate offset inside that plane
  int16_t wslot;     // tap index in the weight / gradient layout
};

// Output pixel geometry: tile-local (n,h,w) -> element offset ((n*H + h*sh+oh)*W + w*sw+ow)*ld
struct PixGeom {
  int H, W, sh, sw, oh, ow;
  long long ld;
  long long sn;      // image stride in elements; 0 -> H*W*ld (dense NHWC)
};

struct TcParams {
  // problem
  int M, N, K;                 // GEMM: per-batch sizes. CONV: N = Cout, K unused. WGRAD: see below
  int num_k_blocks;            // GEMM/CONV: k-blocks of 64 per CTA. WGRAD: total pixel tiles
  int nz0;                     // batch index z = z0 + nz0*z1
  int ksplit;                  // split-K factor (grid.z = batches*ksplit); requires EPI_ATOMIC when > 1
  // conv geometry
  int ntaps, cin_blocks;       // k-block kb -> tap kb / cin_blocks, channel block kb % cin_blocks
  int BW, BH, BNI;             // A box = {64, BW, BH, BNI} pixels (BW*BH*BNI <= 128)
  int tiles_w, tiles_h;        // tiles per image plane; m-tile index -> (nt, ht, wt)
  int NIMG;                    // images (for masking)
  int vH, vW;                  // valid tile-space extent (rows >= vH / cols >= vW are masked)
  PixGeom og, rg;              // output / residual geometry
  TapDesc taps[16];
  // wgrad
  int KK, Cin, Cout;           // dW[co][ci][KK] fp32, atomically accumulated
  int taps_per_cta;            // 1, or 2 when Cin == 64 (two taps stacked along M)
  int pix_per_box;             // BW*BH
  // epilogue
  uint32_t flags;
  float alpha;
  const float* bias;
  const float* scale;
  const float* res_f32;
  const __nv_bfloat16* res_bf16;
  const __nv_bfloat16* aux_bf16;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  long long ldc, ldr;                   // GEMM: row strides of out / residual
  long long osz0, osz1, rsz0, rsz1;     // GEMM: batch strides of out / residual
  int vec_ok;                           // all epilogue pointers / strides allow 16-byte vector access
};

template <int MODE, int A_MN, int B_MN, int BN>
struct TcCfg {
  static constexpr int BM = 128;
  static constexpr int PMAX = 128;                                      // WGRAD: max pixels per stage
  static constexpr int A_BYTES = (MODE == MODE_WGRAD) ? 2 * PMAX * 128 : BM * 128;
  static constexpr int B_BYTES = (MODE == MODE_WGRAD) ? (BN / 64) * PMAX * 128 : BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  static constexpr int THREADS = 192;
  static constexpr int MAX_STAGES = 8;
};

__device__ __forceinline__ long long pix_off(const PixGeom& g, int n, int h, int w) {
  const long long sn = g.sn ? g.sn : static_cast<long long>(g.H) * g.W * g.ld;
  return n * sn + (static_cast<long long>(h * g.sh + g.oh) * g.W + (w * g.sw + g.ow)) * g.ld;
}

template <int MODE, int A_MN, int B_MN, int BN>
__global__ void __launch_bounds__(192, 1)
tc_kernel(const __grid_constant__ CUtensorMap amap0, const __grid_constant__ CUtensorMap amap1,
          const __grid_constant__ CUtensorMap amap2, const __grid_constant__ CUtensorMap amap3,
          const __grid_constant__ CUtensorMap bmap, const __grid_constant__ TcParams p, int num_stages) {
  using Cfg = TcCfg<MODE, A_MN, B_MN, BN>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages * STAGE_BYTES] tiles (1024-aligned), then barriers
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + num_stages * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + Cfg::MAX_STAGES;
  uint64_t* accum_bar = empty_bar + Cfg::MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---------------------------------------------------------------- tile coordinates
  const int n0 = blockIdx.x * BN;
  int z = blockIdx.z, split = 0;
  if (p.ksplit > 1) { split = z % p.ksplit; z = z / p.ksplit; }
  const int z0 = z % p.nz0, z1 = z / p.nz0;
  int m0 = 0, tn = 0, th = 0, tw = 0;         // GEMM row origin / CONV tile origin
  int kb_begin = 0, kb_end = p.num_k_blocks;
  if (MODE == MODE_GEMM) {
    m0 = blockIdx.y * Cfg::BM;
  } else if (MODE == MODE_CONV) {
    int t = blockIdx.y;
    tw = t % p.tiles_w; t /= p.tiles_w;
    th = t % p.tiles_h; tn = t / p.tiles_h;
  }
  if (p.ksplit > 1) {
    int per = (p.num_k_blocks + p.ksplit - 1) / p.ksplit;
    kb_begin = split * per;
    kb_end = min(p.num_k_blocks, kb_begin + per);
  }
  const int nkb = max(0, kb_end - kb_begin);
  // WGRAD: blockIdx.y = (tap group, ci tile); blockIdx.x = co tile
  int wg_tap0 = 0, wg_ci0 = 0;
  if (MODE == MODE_WGRAD) {
    const int ci_tiles = (p.taps_per_cta == 2) ? 1 : (p.Cin / 128);
    wg_ci0 = (blockIdx.y % ci_tiles) * 128;
    wg_tap0 = (blockIdx.y / ci_tiles) * p.taps_per_cta;
  }

  // ---------------------------------------------------------------- setup
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&amap0);
    prefetch_tmap(&bmap);
    for (int s = 0; s < num_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ============================================================== TMA producer
    if (lane == 0) {
      const CUtensorMap* amaps[4] = {&amap0, &amap1, &amap2, &amap3};
      for (int i = 0; i < nkb; ++i) {
        const int kb = kb_begin + i;
        const int s = i % num_stages;
        const uint32_t ph = (i / num_stages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
        uint8_t* sb = sa + Cfg::A_BYTES;
        if (MODE == MODE_GEMM) {
          const uint32_t a_bytes = Cfg::BM * 128;
          const uint32_t b_bytes = BN * 128;
          mbar_expect_tx(&full_bar[s], a_bytes + b_bytes);
          if (A_MN) {
            tma_load_4d(sa, &amap0, &full_bar[s], m0, kb * 64, z0, z1);
            tma_load_4d(sa + 64 * 128, &amap0, &full_bar[s], m0 + 64, kb * 64, z0, z1);
          } else {
            tma_load_4d(sa, &amap0, &full_bar[s], kb * 64, m0, z0, z1);
          }
          if (B_MN) {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)
              tma_load_4d(sb + c * 64 * 128, &bmap, &full_bar[s], n0 + c * 64, kb * 64, z0, z1);
          } else {
            tma_load_4d(sb, &bmap, &full_bar[s], kb * 64, n0, z0, z1);
          }
        } else if (MODE == MODE_CONV) {
          const int tap = kb / p.cin_blocks, cb = kb % p.cin_blocks;
          const TapDesc td = p.taps[tap];
          const uint32_t a_bytes = p.BW * p.BH * p.BNI * 128;
          mbar_expect_tx(&full_bar[s], a_bytes + BN * 128);
          tma_load_4d(sa, amaps[td.map], &full_bar[s], cb * 64, tw * p.BW + td.dw, th * p.BH + td.dh, tn * p.BNI);
          tma_load_2d(sb, &bmap, &full_bar[s], (td.wslot * p.cin_blocks + cb) * 64, n0);
        } else {
          // WGRAD: kb = pixel tile index -> (n, ht, wt)
          int t = kb;
          const int wt = t % p.tiles_w; t /= p.tiles_w;
          const int ht = t % p.tiles_h; const int n = t / p.tiles_h;
          const uint32_t box = p.pix_per_box * 128;
          mbar_expect_tx(&full_bar[s], box * (2 + BN / 64));
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int tap = (p.taps_per_cta == 2) ? min(wg_tap0 + c, p.ntaps - 1) : wg_tap0;
            const int ci = (p.taps_per_cta == 2) ? 0 : wg_ci0 + c * 64;
            const TapDesc td = p.taps[tap];
            tma_load_4d(sa + c * box, amaps[td.map], &full_bar[s], ci, wt * p.BW + td.dw, ht * p.BH + td.dh, n);
          }
#pragma unroll
          for (int c = 0; c < BN / 64; ++c)
            tma_load_4d(sb + c * box, &bmap, &full_bar[s], n0 + c * 64, wt * p.BW, ht * p.BH, n);
        }
      }
    }
  } else if (warp == 1) {
    // ============================================================== MMA issuer (one elected lane)
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(Cfg::BM, BN, A_MN, B_MN);
      const int ksteps = (MODE == MODE_WGRAD) ? (p.pix_per_box / 16) : 4;
      const uint32_t lbo = (MODE == MODE_WGRAD) ? p.pix_per_box * 128 : 64 * 128;
      for (int i = 0; i < nkb; ++i) {
        const int s = i % num_stages;
        const uint32_t ph = (i / num_stages) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES);
        const uint32_t sb = sa + Cfg::A_BYTES;
        for (int k = 0; k < ksteps; ++k) {
          const uint64_t ad = A_MN ? make_smem_desc(sa + k * 2048, lbo, 1024) : make_smem_desc(sa + k * 32, 16, 1024);
          const uint64_t bd = B_MN ? make_smem_desc(sb + k * 2048, lbo, 1024) : make_smem_desc(sb + k * 32, 16, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);          // frees this smem stage when the MMAs above retire
      }
      umma_commit(accum_bar);                // accumulator complete
    }
  } else {
    // ============================================================== epilogue warps (2..5)
    const int q = warp & 3;                  // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;             // accumulator row handled by this thread
    mbar_wait(accum_bar, 0);
    tc_fence_after();

    bool row_ok = false;
    long long o_off = 0, r_off = 0;
    if (MODE == MODE_GEMM) {
      const int gm = m0 + r;
      row_ok = gm < p.M;
      o_off = z0 * p.osz0 + z1 * p.osz1 + static_cast<long long>(gm) * p.ldc;
      r_off = z0 * p.rsz0 + z1 * p.rsz1 + static_cast<long long>(gm) * p.ldr;
    } else if (MODE == MODE_CONV) {
      const int w = r % p.BW, h = (r / p.BW) % p.BH, ni = r / (p.BW * p.BH);
      const int gw = tw * p.BW + w, gh = th * p.BH + h, gn = tn * p.BNI + ni;
      row_ok = (ni < p.BNI) && gw < p.vW && gh < p.vH && gn < p.NIMG;
      o_off = pix_off(p.og, gn, gh, gw);
      r_off = pix_off(p.rg, gn, gh, gw);
    } else {
      // WGRAD: row -> (tap, ci)
      int tap, ci;
      if (p.taps_per_cta == 2) { tap = wg_tap0 + (r >> 6); ci = r & 63; }
      else { tap = wg_tap0; ci = wg_ci0 + r; }
      row_ok = tap < p.ntaps && ci < p.Cin;
      o_off = static_cast<long long>(ci) * p.KK + p.taps[min(tap, p.ntaps - 1)].wslot;
    }
    if (nkb == 0) row_ok = false;            // nothing accumulated (empty split): TMEM is garbage

#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      uint32_t v[32];
      __syncwarp();
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (!row_ok) continue;
      const int gn0 = n0 + c0;
      if (MODE == MODE_WGRAD) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int co = gn0 + j;
          if (co < p.Cout) {
            const float sc = (p.flags & EPI_SCALE) ? __ldg(p.scale + co) * p.alpha : p.alpha;
            atomicAdd(p.out_f32 + static_cast<long long>(co) * p.Cin * p.KK + o_off, __uint_as_float(v[j]) * sc);
          }
        }
        continue;
      }
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
      const bool full = p.vec_ok && (gn0 + 32 <= p.N);
      if (p.flags & EPI_SCALE) {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (full || gn0 + j < p.N) f[j] *= __ldg(p.scale + gn0 + j);
      }
      if (p.flags & EPI_BIAS) {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (full || gn0 + j < p.N) f[j] += __ldg(p.bias + gn0 + j);
      }
      if (p.flags & EPI_RES_F32) {
        const float* rp = p.res_f32 + r_off + gn0;
        if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 t = *reinterpret_cast<const float4*>(rp + j);
            f[j] += t.x; f[j + 1] += t.y; f[j + 2] += t.z; f[j + 3] += t.w;
          }
        } else {
          for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) f[j] += rp[j];
        }
      }
      if (p.flags & EPI_RES_BF16) {
        const __nv_bfloat16* rp = p.res_bf16 + r_off + gn0;
        if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const uint4 t = *reinterpret_cast<const uint4*>(rp + j);
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float2 x = __bfloat1622float2(h2[e]); f[j + 2 * e] += x.x; f[j + 2 * e + 1] += x.y; }
          }
        } else {
          for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) f[j] += __bfloat162float(rp[j]);
        }
      }
      if (p.flags & EPI_RELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      if (p.flags & EPI_MASK_POS) {
        const __nv_bfloat16* ap = p.aux_bf16 + o_off + gn0;
        if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const uint4 t = *reinterpret_cast<const uint4*>(ap + j);
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 x = __bfloat1622float2(h2[e]);
              if (!(x.x > 0.f)) f[j + 2 * e] = 0.f;
              if (!(x.y > 0.f)) f[j + 2 * e + 1] = 0.f;
            }
          }
        } else {
          for (int j = 0; j < 32; ++j) if (gn0 + j < p.N && !(__bfloat162float(ap[j]) > 0.f)) f[j] = 0.f;
        }
      }
      if (p.flags & EPI_OUT_F32) {
        float* op = p.out_f32 + o_off + gn0;
        if (p.flags & EPI_ATOMIC) {
#pragma unroll
          for (int j = 0; j < 32; ++j) if (full || gn0 + j < p.N) atomicAdd(op + j, f[j]);
        } else if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(op + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
        } else {
          for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) op[j] = f[j];
        }
      }
      if (p.flags & EPI_OUT_BF16) {
        __nv_bfloat16* op = p.out_bf16 + o_off + gn0;
        if (full) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 t;
            __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
            for (int e = 0; e < 4; ++e) h2[e] = __floats2bfloat162_rn(f[j + 2 * e], f[j + 2 * e + 1]);
            *reinterpret_cast<uint4*>(op + j) = t;
          }
        } else {
          for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) op[j] = __float2bfloat16_rn(f[j]);
        }
      }
    }
  }

  // ---------------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace act
