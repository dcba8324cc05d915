// One warp-specialised tcgen05 kernel template shared by every dense contraction on the ACT path:
//   MODE_GEMM  : batched D[z][M,N] = A[z][M,K] * B[z][N,K]^T, each operand K-major or MN-major
//   MODE_CONV  : implicit-GEMM convolution / transposed convolution over NHWC bf16 (tap table, TMA boxes)
//   MODE_WGRAD : convolution weight gradient, reduction over pixels, both operands MN-major, split-K atomics
// The kernel is PERSISTENT: one CTA per SM walks the tile list (tile = blockIdx.x + i*gridDim.x) so that the TMA
// producer runs ahead across tile boundaries (the smem ring never drains) and the epilogue of tile i overlaps the
// MMAs of tile i+1 through two TMEM accumulator buffers.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> global).
// Operands land in 128B-swizzled shared memory via TMA; accumulators live in TMEM (fp32).
#pragma once
#include "ptx.cuh"
#include "dropout.cuh"

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
  EPI_DROPOUT = 1u << 10,    // train-mode dropout of (acc + bias) before the residual is added (MODE_GEMM, un-batched)
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
  int ksplit;                  // split-K factor (gz = batches*ksplit); requires EPI_ATOMIC when > 1
  int b_res;                   // MODE_CONV: 1 = the whole weight operand (num_k_blocks x [BN x 64]) is loaded ONCE per CTA and stays in smem
  uint32_t skip_mask;          // conv3_kernel: bit (tap * cin_blocks + cb) set -> that weight block is all zero, skip it
  int gx, gy, gz;              // tile grid: x = N tiles (fastest), y = M tiles / pixel tiles / wgrad row groups, z = batch*split
  // conv geometry
  int ntaps, cin_blocks;       // k-block kb -> tap kb / cin_blocks, channel block kb % cin_blocks
  int BW, BH, BNI;             // A box = {64, BW, BH, BNI} pixels (BW*BH*BNI <= 128)
  int PW;                      // conv3 (halo-resident) mode: patch row pitch = BW + 2
  int tiles_w, tiles_h;        // tiles per image plane; m-tile index -> (nt, ht, wt)
  int NIMG;                    // images (for masking)
  int vH, vW;                  // valid tile-space extent (rows >= vH / cols >= vW are masked)
  PixGeom og, rg;              // output / residual geometry
  TapDesc taps[16];
  // wgrad
  int KK, Cin, Cout;           // staged dW: T[KK][Cout][Cin] fp32, atomically accumulated (line-coalesced reductions)
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
  int vec32_ok;                         // ... and the outputs allow 32-byte (sector-wide) stores
  int debug;                            // experiment switches (ACT_TC_DEBUG): 1 = skip epilogue math+stores, 2 = skip TMA loads
  DropDesc drop;                        // EPI_DROPOUT: element index = row * drop_ld + column
  long long drop_ld;
  int pdl_late;                         // 1: let the dependent kernel start when this CTA has issued its last MMA (not at entry)
  unsigned long long* ts;               // diagnostics (act_tc_debug_timestamps): [0] = launch counter, then 8 stamps per launch
};

__device__ __forceinline__ unsigned long long tc_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

template <int MODE, int A_MN, int B_MN, int BN>
struct TcCfg {
  static constexpr int BM = 128;
  static constexpr int PMAX = 128;                                      // WGRAD: max pixels per stage
  static constexpr int A_BYTES = (MODE == MODE_WGRAD) ? 2 * PMAX * 128 : BM * 128;
  static constexpr int B_BYTES = (MODE == MODE_WGRAD) ? (BN / 64) * PMAX * 128 : BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int ACC_COLS = BN;                                   // one accumulator buffer
  static constexpr int TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;
  static constexpr int THREADS = 320;      // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue (two per TMEM lane quadrant)
  static constexpr int MAX_STAGES = 8;
};

__device__ __forceinline__ long long pix_off(const PixGeom& g, int n, int h, int w) {
  const long long sn = g.sn ? g.sn : static_cast<long long>(g.H) * g.W * g.ld;
  return n * sn + (static_cast<long long>(h * g.sh + g.oh) * g.W + (w * g.sw + g.ow)) * g.ld;
}

// Per-tile coordinates, recomputed by each warp role from the linear tile index.
struct TileCoord {
  int n0, z0, z1, m0, tn, th, tw, kb_begin, nkb, wg_tap0, wg_ci0;
};

template <int MODE, int BN>
__device__ __forceinline__ TileCoord decode_tile(const TcParams& p, int t) {
  TileCoord c;
  const int bx = t % p.gx; t /= p.gx;
  const int by = t % p.gy;
  int z = t / p.gy;
  c.n0 = bx * BN;
  int split = 0;
  if (p.ksplit > 1) { split = z % p.ksplit; z = z / p.ksplit; }
  c.z0 = z % p.nz0; c.z1 = z / p.nz0;
  c.m0 = 0; c.tn = 0; c.th = 0; c.tw = 0; c.wg_tap0 = 0; c.wg_ci0 = 0;
  if (MODE == MODE_GEMM) {
    c.m0 = by * 128;
  } else if (MODE == MODE_CONV) {
    int u = by;
    c.tw = u % p.tiles_w; u /= p.tiles_w;
    c.th = u % p.tiles_h; c.tn = u / p.tiles_h;
  } else {
    const int ci_tiles = (p.taps_per_cta == 2) ? 1 : (p.Cin / 128);
    c.wg_ci0 = (by % ci_tiles) * 128;
    c.wg_tap0 = (by / ci_tiles) * p.taps_per_cta;
  }
  int kb_begin = 0, kb_end = p.num_k_blocks;
  if (p.ksplit > 1) {
    const int per = (p.num_k_blocks + p.ksplit - 1) / p.ksplit;
    kb_begin = split * per;
    kb_end = min(p.num_k_blocks, kb_begin + per);
  }
  c.kb_begin = kb_begin;
  c.nkb = max(0, kb_end - kb_begin);
  return c;
}

__device__ __forceinline__ void ldg32(const float* __restrict__ ptr, float (&dst)[32]) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(ptr + j));
    dst[j] = t.x; dst[j + 1] = t.y; dst[j + 2] = t.z; dst[j + 3] = t.w;
  }
}

// Fused epilogue of one 32-column accumulator chunk of one row: alpha, per-column scale / bias, fp32 / bf16 residual,
// ReLU, ReLU-mask against a saved activation, fp32 / bf16 / atomic output.  o_off / r_off = element offsets of the row.
// `addends` = false for k-splits other than the first: bias and residual are added exactly once per output element.
__device__ __forceinline__ void epi_store_chunk(const TcParams& p, const uint32_t (&v)[32], int gn0, long long o_off,
                                                long long r_off, bool addends = true, int drop_row = 0) {
  float f[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * p.alpha;
  const bool full = p.vec_ok && (gn0 + 32 <= p.N);
  if (p.flags & EPI_SCALE) {
    if (full) {
      float sc[32];
      ldg32(p.scale + gn0, sc);
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] *= sc[j];
    } else {
      for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) f[j] *= __ldg(p.scale + gn0 + j);
    }
  }
  if ((p.flags & EPI_BIAS) && addends) {
    if (full) {
      float bi[32];
      ldg32(p.bias + gn0, bi);
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] += bi[j];
    } else {
      for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) f[j] += __ldg(p.bias + gn0 + j);
    }
  }
  if (p.flags & EPI_DROPOUT)               // dropout(linear(x)) -- commutes with the ReLU below (scale > 0, mask in {0, 1})
    drop_apply32(p.drop, static_cast<unsigned long long>(drop_row) * p.drop_ld + gn0, __ldg(p.drop.step), f);
  if ((p.flags & EPI_RES_F32) && addends) {
    const float* rp = p.res_f32 + r_off + gn0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 u = *reinterpret_cast<const float4*>(rp + j);
        f[j] += u.x; f[j + 1] += u.y; f[j + 2] += u.z; f[j + 3] += u.w;
      }
    } else {
      for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) f[j] += rp[j];
    }
  }
  if ((p.flags & EPI_RES_BF16) && addends) {
    const __nv_bfloat16* rp = p.res_bf16 + r_off + gn0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        const uint4 u = *reinterpret_cast<const uint4*>(rp + j);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
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
        const uint4 u = *reinterpret_cast<const uint4*>(ap + j);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
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
      if (full) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) red_add_v4(op + j, f[j], f[j + 1], f[j + 2], f[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) atomicAdd(op + j, f[j]);
      }
    } else if (full && p.vec32_ok) {
#pragma unroll
      for (int j = 0; j < 32; j += 8)
        st_global_v8(op + j, __float_as_uint(f[j]), __float_as_uint(f[j + 1]), __float_as_uint(f[j + 2]), __float_as_uint(f[j + 3]),
                     __float_as_uint(f[j + 4]), __float_as_uint(f[j + 5]), __float_as_uint(f[j + 6]), __float_as_uint(f[j + 7]));
    } else if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(op + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
    } else {
      for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) op[j] = f[j];
    }
  }
  if (p.flags & EPI_OUT_BF16) {
    __nv_bfloat16* op = p.out_bf16 + o_off + gn0;
    if (full && p.vec32_ok) {
      uint32_t w[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * e], f[2 * e + 1]);
        w[e] = *reinterpret_cast<const uint32_t*>(&h);
      }
      st_global_v8(op, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
      st_global_v8(op + 16, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]);
    } else if (full) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u;
        __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) h2[e] = __floats2bfloat162_rn(f[j + 2 * e], f[j + 2 * e + 1]);
        *reinterpret_cast<uint4*>(op + j) = u;
      }
    } else {
      for (int j = 0; j < 32; ++j) if (gn0 + j < p.N) op[j] = __float2bfloat16_rn(f[j]);
    }
  }
}

template <int MODE, int A_MN, int B_MN, int BN>
__global__ void __launch_bounds__(320, 1)
tc_kernel(const __grid_constant__ CUtensorMap amap0, const __grid_constant__ CUtensorMap amap1,
          const __grid_constant__ CUtensorMap amap2, const __grid_constant__ CUtensorMap amap3,
          const __grid_constant__ CUtensorMap bmap, const __grid_constant__ TcParams p, int num_stages) {
  using Cfg = TcCfg<MODE, A_MN, B_MN, BN>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages * STAGE_BYTES] tiles (1024-aligned), then barriers
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // resident weights (MODE_CONV, p.b_res): stages hold the A tile only, the weight operand follows the ring and is
  // loaded once per CTA -- a persistent CTA re-reads the same [Cout x K] weights for every tile otherwise, and for the
  // stem that re-streaming was half of the L2->SM traffic that bounds it
  const bool b_res = (MODE == MODE_CONV) && p.b_res != 0;
  const uint32_t stage_bytes = b_res ? Cfg::A_BYTES : Cfg::STAGE_BYTES;
  const uint32_t res_bytes = b_res ? static_cast<uint32_t>(p.num_k_blocks) * BN * 128u : 0u;
  uint8_t* bres = smem + num_stages * stage_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bres + res_bytes);
  uint64_t* empty_bar = full_bar + Cfg::MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + Cfg::MAX_STAGES;      // [2] accumulator buffer complete (MMA -> epilogue)
  uint64_t* tempty_bar = tfull_bar + 2;                   // [2] accumulator buffer drained (epilogue -> MMA)
  uint64_t* bres_bar = tempty_bar + 2;                    // resident weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.gx * p.gy * p.gz;

  // ---------------------------------------------------------------- setup (overlaps the previous kernel's tail)
  // stage timestamps of CTA 0 (diagnostics only, p.ts == nullptr in production): slot claimed by thread 0
  unsigned long long*& ts_slot = *reinterpret_cast<unsigned long long**>(tmem_slot + 2);   // (inside the 16 bytes behind the barriers)
  if (p.ts != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    ts_slot = p.ts + 1 + 8 * atomicAdd(p.ts, 1ull);
    ts_slot[0] = tc_globaltimer();
  }
  const bool stamp = p.ts != nullptr && blockIdx.x == 0;
  // Programmatic dependent launch.  Triggering at entry lets the next kernel's CTAs become resident at once -- but they then
  // sit in griddepcontrol.wait holding a whole SM (ring + TMEM) for as long as this kernel runs (measured: CTAs resident
  // 9-11 us for a 4.3 us launch period), which starves the other branches of a concurrent step.  The late trigger (after
  // this CTA's last MMA has been issued) still hides the dependent's set-up under our last epilogue.
  if (!p.pdl_late) pdl_trigger();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&amap0);
    prefetch_tmap(&bmap);
    for (int s = 0; s < num_stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], (blockDim.x >> 5) - 2); }
    mbar_init(bres_bar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (stamp && threadIdx.x == 0) ts_slot[1] = tc_globaltimer();
  pdl_wait();                                  // operands / residuals written by the previous kernel are visible from here
  if (stamp && threadIdx.x == 0) ts_slot[2] = tc_globaltimer();

  if (warp == 0) {
    // ============================================================== TMA producer
    if (elect_one()) {
      const CUtensorMap* amaps[4] = {&amap0, &amap1, &amap2, &amap3};
      uint32_t s = 0, ph = 0;                           // ring position / phase, carried across tiles (no div/mod in the loop)
      if (MODE == MODE_CONV && b_res) {
        mbar_expect_tx(bres_bar, res_bytes);
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          const int tap = kb / p.cin_blocks, cb = kb % p.cin_blocks;
          tma_load_2d(bres + static_cast<uint32_t>(kb) * BN * 128u, &bmap, bres_bar, (p.taps[tap].wslot * p.cin_blocks + cb) * 64, 0);
        }
      }
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord c = decode_tile<MODE, BN>(p, t);
        for (int i = 0; i < c.nkb; ++i) {
          const int kb = c.kb_begin + i;
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * stage_bytes;
          uint8_t* sb = sa + Cfg::A_BYTES;
          uint64_t* const fb = &full_bar[s];
          if (++s == static_cast<uint32_t>(num_stages)) { s = 0; ph ^= 1; }
          if (p.debug & 2) { mbar_arrive(fb); continue; }
          if (MODE == MODE_GEMM) {
            const uint32_t a_bytes = Cfg::BM * 128;
            const uint32_t b_bytes = BN * 128;
            mbar_expect_tx(fb, a_bytes + b_bytes);
            if (A_MN) {
              tma_load_4d(sa, &amap0, fb, c.m0, kb * 64, c.z0, c.z1);
              tma_load_4d(sa + 64 * 128, &amap0, fb, c.m0 + 64, kb * 64, c.z0, c.z1);
            } else {
              tma_load_4d(sa, &amap0, fb, kb * 64, c.m0, c.z0, c.z1);
            }
            if (B_MN) {
#pragma unroll
              for (int q = 0; q < BN / 64; ++q)
                tma_load_4d(sb + q * 64 * 128, &bmap, fb, c.n0 + q * 64, kb * 64, c.z0, c.z1);
            } else {
              tma_load_4d(sb, &bmap, fb, kb * 64, c.n0, c.z0, c.z1);
            }
          } else if (MODE == MODE_CONV) {
            const int tap = kb / p.cin_blocks, cb = kb % p.cin_blocks;
            const TapDesc td = p.taps[tap];
            const uint32_t a_bytes = p.BW * p.BH * p.BNI * 128;
            mbar_expect_tx(fb, a_bytes + (b_res ? 0u : BN * 128u));
            tma_load_4d(sa, amaps[td.map], fb, cb * 64, c.tw * p.BW + td.dw, c.th * p.BH + td.dh, c.tn * p.BNI);
            if (!b_res) tma_load_2d(sb, &bmap, fb, (td.wslot * p.cin_blocks + cb) * 64, c.n0);
          } else {
            // WGRAD: kb = pixel tile index -> (n, ht, wt)
            int u = kb;
            const int wt = u % p.tiles_w; u /= p.tiles_w;
            const int ht = u % p.tiles_h; const int n = u / p.tiles_h;
            const uint32_t box = p.pix_per_box * 128;
            mbar_expect_tx(fb, box * (2 + BN / 64));
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              const int tap = (p.taps_per_cta == 2) ? min(c.wg_tap0 + q, p.ntaps - 1) : c.wg_tap0;
              const int ci = (p.taps_per_cta == 2) ? 0 : c.wg_ci0 + q * 64;
              const TapDesc td = p.taps[tap];
              tma_load_4d(sa + q * box, amaps[td.map], fb, ci, wt * p.BW + td.dw, ht * p.BH + td.dh, n);
            }
#pragma unroll
            for (int q = 0; q < BN / 64; ++q)
              tma_load_4d(sb + q * box, &bmap, fb, c.n0 + q * 64, wt * p.BW, ht * p.BH, n);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================================================== MMA issuer (one elected lane)
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(Cfg::BM, BN, A_MN, B_MN);
      const int ksteps = (MODE == MODE_WGRAD) ? (p.pix_per_box / 16) : 4;
      const uint32_t lbo = (MODE == MODE_WGRAD) ? p.pix_per_box * 128 : 64 * 128;
      const uint32_t sbase = smem_u32(smem);
      const uint64_t adesc0 = A_MN ? make_smem_desc(sbase, lbo, 1024) : make_smem_desc(sbase, 16, 1024);
      const uint64_t bdesc0 = B_MN ? make_smem_desc(sbase + Cfg::A_BYTES, lbo, 1024) : make_smem_desc(sbase + Cfg::A_BYTES, 16, 1024);
      uint32_t s = 0, ph = 0, lt = 0;                   // ring position / phase carried across tiles
      const uint32_t stage16 = stage_bytes >> 4;
      const uint64_t bres_desc0 = make_smem_desc(smem_u32(bres), 16, 1024);
      if (MODE == MODE_CONV && b_res) { mbar_wait(bres_bar, 0); tc_fence_after(); }
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
        const TileCoord c = decode_tile<MODE, BN>(p, t);
        const uint32_t acc = lt & 1;
        mbar_wait(&tempty_bar[acc], ((lt >> 1) & 1) ^ 1);      // epilogue drained this buffer (first use: immediate)
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * Cfg::ACC_COLS;
        for (int i = 0; i < c.nkb; ++i) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (stamp && lt == 0 && i == 0) ts_slot[3] = tc_globaltimer();
          // descriptors differ from the stage-0 / k-0 ones only in the 14-bit start-address field (16-byte units)
          const uint32_t soff = s * stage16;
          uint64_t ad = adesc0 + soff, bd = bdesc0 + soff;
          if (MODE == MODE_CONV && b_res) bd = bres_desc0 + ((static_cast<uint32_t>(c.kb_begin + i) * BN * 128u) >> 4);
          if (MODE == MODE_WGRAD) {
            for (int k = 0; k < ksteps; ++k, ad += 128, bd += 128) umma_bf16(tmem_d, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k, ad += (A_MN ? 128 : 2), bd += (B_MN ? 128 : 2))
              umma_bf16(tmem_d, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);          // frees this smem stage when the MMAs above retire
          if (++s == static_cast<uint32_t>(num_stages)) { s = 0; ph ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);          // accumulator complete
        if (stamp) ts_slot[4] = tc_globaltimer();
      }
      if (p.pdl_late) pdl_trigger();
    }
  } else {
    // ============================================================== epilogue warps (2..9)
    // Two warps per TMEM lane quadrant, each draining half of the tile's columns: for the single-round launches that
    // dominate the step the epilogue is exposed latency, so its width matters more than its instruction count.
    const int q = warp & 3;                  // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;             // accumulator row handled by this thread
    // 8 epilogue warps (320 threads): two per quadrant, half of the columns each; 4 epilogue warps (192 threads, the
    // SM-sharing launch form: two CTAs per SM must fit the register file): one per quadrant, all columns
    const int halves = ((blockDim.x >> 5) - 2) >> 2;
    const int cbeg = ((warp - 2) >> 2) * (BN / halves), cend = cbeg + BN / halves;
    uint32_t lt = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
      const TileCoord c = decode_tile<MODE, BN>(p, t);
      const uint32_t acc = lt & 1;
      const int n0 = c.n0;

      bool row_ok = false;
      long long o_off = 0, r_off = 0;
      if (MODE == MODE_GEMM) {
        const int gm = c.m0 + r;
        row_ok = gm < p.M;
        o_off = c.z0 * p.osz0 + c.z1 * p.osz1 + static_cast<long long>(gm) * p.ldc;
        r_off = c.z0 * p.rsz0 + c.z1 * p.rsz1 + static_cast<long long>(gm) * p.ldr;
      } else if (MODE == MODE_CONV) {
        const int w = r % p.BW, h = (r / p.BW) % p.BH, ni = r / (p.BW * p.BH);
        const int gw = c.tw * p.BW + w, gh = c.th * p.BH + h, gn = c.tn * p.BNI + ni;
        row_ok = (ni < p.BNI) && gw < p.vW && gh < p.vH && gn < p.NIMG;
        o_off = pix_off(p.og, gn, gh, gw);
        r_off = pix_off(p.rg, gn, gh, gw);
      } else {
        // WGRAD: row -> (tap, ci)
        int tap, ci;
        if (p.taps_per_cta == 2) { tap = c.wg_tap0 + (r >> 6); ci = r & 63; }
        else { tap = c.wg_tap0; ci = c.wg_ci0 + r; }
        row_ok = tap < p.ntaps && ci < p.Cin;
        o_off = static_cast<long long>(p.taps[min(tap, p.ntaps - 1)].wslot) * p.Cout * p.Cin + ci;
      }
      if (c.nkb == 0) row_ok = false;          // nothing accumulated (empty split): TMEM is garbage

      mbar_wait(&tfull_bar[acc], (lt >> 1) & 1);
      tc_fence_after();
      if (stamp && warp == 2 && lane == 0) ts_slot[5] = tc_globaltimer();
      const uint32_t tmem_acc = tmem_base + acc * Cfg::ACC_COLS + (static_cast<uint32_t>(q * 32) << 16);

#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32(tmem_acc + c0, v);
        tmem_ld_wait();
        if (c0 + 32 >= cend) {                 // last chunk is in registers: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        if (!row_ok || (p.debug & 1)) continue;
        const int gn0 = n0 + c0;
        if (MODE == MODE_WGRAD) {
          // staged layout T[wslot][co][ci]: the 32 lanes of a warp hold 32 consecutive ci, so every reduction
          // instruction covers one contiguous 128-byte line; the caller folds T into dW[co][ci][kk]
          float* op = p.out_f32 + o_off + static_cast<long long>(gn0) * p.Cin;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (gn0 + j < p.Cout) atomicAdd(op + static_cast<long long>(j) * p.Cin, __uint_as_float(v[j]) * p.alpha);
          continue;
        }
        epi_store_chunk(p, v, gn0, o_off, r_off, p.ksplit <= 1 || c.kb_begin == 0, c.m0 + r);
      }
    }
  }

  // ---------------------------------------------------------------- teardown
  if (stamp && warp == 2 && lane == 0) ts_slot[6] = tc_globaltimer();
  tc_fence_before();
  __syncthreads();
  if (stamp && threadIdx.x == 0) ts_slot[7] = tc_globaltimer();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace act
