// Attention scores fused with their row-wise softmax (forward) or softmax backward, for head_dim 64 and up to 320 keys:
//   FWD:  P  = softmax(scale * Q K^T + key_padding_mask)                      (transformer.py:194,254-255 via nn.MHA)
//   BWD:  dS = scale * P o (dP - rowsum(dP o P)),   dP = dO V^T               (its autograd)
// One CTA tile = 128 query rows x ALL keys of one (batch, head): the [128 x W] fp32 score tile lives in TMEM
// (W = 128 or 320 columns), so the row reductions never leave the SM and the fp32 score matrix never touches HBM
// (the unfused path writes S fp32, reads it back in a softmax kernel and writes P: 3 x the traffic and one more launch).
// Warp roles as in tc_kernel: warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue -- two warps per TMEM lane
// quadrant, each owning half of the row's columns in registers; the two halves of a row exchange max / sum through
// shared memory (named barrier over the 256 epilogue threads).
#pragma once
#include "tc_kernel.cuh"
#include "dropout.cuh"

namespace act {

struct AttnParams {
  int Lq, Lk, Lkp;              // query rows, keys, padded row stride of P / dS (multiple of 8)
  int heads, batch, qtiles;
  float scale;
  const uint8_t* mask;          // [batch, Lk] key padding mask (1 = masked) or null        (FWD)
  const __nv_bfloat16* P;       // [batch, heads, Lq, Lkp] probabilities                    (BWD input)
  __nv_bfloat16* out;           // P (FWD) or dS (BWD), same layout
  int vec32;                    // rows of `out` are 32-byte aligned (Lkp % 16 == 0): whole-sector stores
  // train-mode dropout on the attention weights (nn.MultiheadAttention(dropout=p), transformer.py:194,254-255):
  // FWD also writes out2 = dropout(P) (the operand of P.V; P itself is kept for the softmax backward);
  // BWD applies the same mask to dP = dO V^T before the softmax backward.  Element index = offset in `out`.
  DropDesc drop;
  __nv_bfloat16* out2;
  // chunked kernel (Lk > 320): work item = (batch, head, group of consecutive query tiles); K/V stays resident per item
  int nch, groups, tiles_per_group;
  // BWD, optional: the forward's attention output O = dropout(P) V and the incoming gradient dO as plain row views (row
  // stride ld, head stride sz0, batch stride sz1, elements).  rowsum(P o dP) = dO_i . O_i (also under dropout), so the
  // chunked kernel needs ONE pass over the keys instead of two when O is given.
  const __nv_bfloat16* O; const __nv_bfloat16* dO;
  long long o_ld, o_sz0, o_sz1, do_ld, do_sz0, do_sz1;
};

// bf16 store of one 32-column chunk of a row (columns [col0, col0+32) clipped to Lkp, Lkp % 8 == 0)
__device__ __forceinline__ void attn_store_chunk(__nv_bfloat16* orow, int col0, const float (&v)[32], int Lkp, int vec32) {
  if (col0 >= Lkp) return;
  uint32_t w[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
    w[e] = *reinterpret_cast<const uint32_t*>(&h);
  }
  if (vec32 && col0 + 32 <= Lkp) {
    st_global_v8(orow + col0, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
    st_global_v8(orow + col0 + 16, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)                       // Lkp is a multiple of 8: whole 16-byte pieces
      if (col0 + e * 8 < Lkp)
        *reinterpret_cast<uint4*>(orow + col0 + e * 8) = make_uint4(w[4 * e], w[4 * e + 1], w[4 * e + 2], w[4 * e + 3]);
  }
}

template <int W>
struct AttnCfg {
  static constexpr int NMMA = W > 256 ? 2 : 1;          // tcgen05 N <= 256: 320 keys = 2 x 160
  static constexpr int NPER = W / NMMA;
  static constexpr int A_BYTES = 128 * 128;             // 128 rows x 64 bf16
  static constexpr int B_BYTES = W * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = 2;
  static constexpr int TMEM_COLS = W <= 128 ? 128 : 512;
  static constexpr int NC = W / 64;                     // 32-column chunks per epilogue warp (half a row)
  static constexpr int SMEM = 1024 + STAGES * STAGE_BYTES + 2 * 2 * 128 * 4 + 4 * 8 + 16;
};

template <int W, int BWD>
__global__ void __launch_bounds__(256, 1)
attn_scores_kernel(const __grid_constant__ CUtensorMap amap, const __grid_constant__ CUtensorMap bmap,
                   const __grid_constant__ AttnParams p) {
  using Cfg = AttnCfg<W>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* xch = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);      // [2 passes][2 halves][128 rows]
  uint64_t* full = reinterpret_cast<uint64_t*>(xch + 2 * 2 * 128);
  uint64_t* tfull = full + Cfg::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);

  // Eight warps, all of them epilogue (two per TMEM lane quadrant; 10 warps would cap the kernel at 168 registers per
  // thread -- three warps on one SM sub-partition -- and spill the 160-column row half).  Thread 0 is also the TMA
  // producer and the MMA issuer: one tile is a single 64-deep k-block (8 MMAs), and with a [128 x 320] fp32 tile the
  // TMEM holds one accumulator, so MMA(i+1) can only start once tile i is in registers anyway.
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool leader = threadIdx.x == 0;
  const int total_tiles = p.qtiles * p.heads * p.batch;
  constexpr uint32_t idesc = make_idesc_bf16(128, Cfg::NPER, 0, 0);

  pdl_trigger();
  if (leader) {
    prefetch_tmap(&amap);
    prefetch_tmap(&bmap);
    for (int s = 0; s < Cfg::STAGES; ++s) mbar_init(&full[s], 1);
    mbar_init(tfull, 1);
    fence_mbar_init();
  }
  if (warp == 0) { __syncwarp(); tmem_alloc(tmem_slot, Cfg::TMEM_COLS); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  auto load_tile = [&](int t, int s) {                     // leader only
    const int qt = t % p.qtiles, hd = (t / p.qtiles) % p.heads, b = t / (p.qtiles * p.heads);
    mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
    uint8_t* st = smem + s * Cfg::STAGE_BYTES;
    tma_load_4d(st, &amap, &full[s], 0, qt * 128, hd, b);                          // rows >= Lq are zero-filled
#pragma unroll
    for (int h = 0; h < Cfg::NMMA; ++h)                                              // keys >= Lk are zero-filled
      tma_load_4d(st + Cfg::A_BYTES + h * Cfg::NPER * 128, &bmap, &full[s], 0, h * Cfg::NPER, hd, b);
  };
  auto mma_tile = [&](uint32_t i) {                        // leader only: tile i of this CTA sits in stage i % STAGES
    const int s = i % Cfg::STAGES;
    mbar_wait(&full[s], (i / Cfg::STAGES) & 1);
    tc_fence_after();
    const uint32_t a0 = smem_u32(smem) + s * Cfg::STAGE_BYTES;
#pragma unroll
    for (int h = 0; h < Cfg::NMMA; ++h) {
      uint64_t ad = make_smem_desc(a0, 16, 1024);
      uint64_t bd = make_smem_desc(a0 + Cfg::A_BYTES + h * Cfg::NPER * 128, 16, 1024);
#pragma unroll
      for (int k = 0; k < 4; ++k, ad += 2, bd += 2) umma_bf16(tmem_base + h * Cfg::NPER, ad, bd, idesc, k > 0 ? 1u : 0u);
    }
    umma_commit(tfull);
  };
  if (leader && static_cast<int>(blockIdx.x) < total_tiles) {
    load_tile(blockIdx.x, 0);
    if (static_cast<int>(blockIdx.x + gridDim.x) < total_tiles) load_tile(blockIdx.x + gridDim.x, 1);
    mma_tile(0);
  }
  __syncwarp();

  {
    const int q = warp & 3, half = warp >> 2;
    const int r = q * 32 + lane;
    const int cbeg = half * (W / 2);
    uint32_t it = 0;
    // leader, once every warp holds tile `it` in registers: refill the stage tile `it` used, then issue tile it+1
    auto next_tile = [&](int t, uint32_t i) {
      const int t1 = t + gridDim.x, t2 = t + 2 * gridDim.x;
      if (t1 >= total_tiles) return;
      tc_fence_after();
      if (t2 < total_tiles) load_tile(t2, i % Cfg::STAGES);
      mma_tile(i + 1);
    };
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
      const int qt = t % p.qtiles, hd = (t / p.qtiles) % p.heads, b = t / (p.qtiles * p.heads);
      const int gm = qt * 128 + r;
      const bool row_ok = gm < p.Lq;
      const long long row_off = ((static_cast<long long>(b) * p.heads + hd) * p.Lq + (row_ok ? gm : 0)) * p.Lkp;
      mbar_wait(tfull, it & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + cbeg;
      float v[Cfg::NC][32];
#pragma unroll
      for (int c = 0; c < Cfg::NC; ++c) {
        uint32_t u[32];
        tmem_ld_32x32(tacc + c * 32, u);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) v[c][j] = __uint_as_float(u[j]);
      }
      tc_fence_before();                                    // the barrier below also tells the leader TMEM is drained

      float* x0 = xch + (half * 128 + r);                   // my slot / my partner's slot, pass 0
      float* y0 = xch + ((half ^ 1) * 128 + r);
      if (!BWD) {
        const uint8_t* mrow = p.mask ? p.mask + static_cast<long long>(b) * p.Lk : nullptr;
        // eight independent partial maxima / sums: with two warps per scheduler a single 160-long dependent chain of
        // FMNMX / FADD is pure latency (ncu: 18 % issue-active before this change)
        float m8[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) m8[e] = -INFINITY;
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c) {
          const int col0 = cbeg + c * 32;
          if (!mrow && col0 + 32 <= p.Lk) {                   // chunk entirely inside the row, no mask: scale only
#pragma unroll
            for (int j = 0; j < 32; ++j) { v[c][j] *= p.scale; m8[j & 7] = fmaxf(m8[j & 7], v[c][j]); }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int col = col0 + j;
              const bool dead = col >= p.Lk || (mrow && __ldg(mrow + min(col, p.Lk - 1)));
              v[c][j] = dead ? -INFINITY : v[c][j] * p.scale;
              m8[j & 7] = fmaxf(m8[j & 7], v[c][j]);
            }
          }
        }
        float mx = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])), fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
        *x0 = mx;
        __syncthreads();
        if (leader) next_tile(t, it);
        __syncwarp();
        mx = fmaxf(mx, *y0);
        if (mx == -INFINITY) mx = 0.f;
        float s8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c)
#pragma unroll
          for (int j = 0; j < 32; ++j) { v[c][j] = __expf(v[c][j] - mx); s8[j & 7] += v[c][j]; }
        float sum = ((s8[0] + s8[1]) + (s8[2] + s8[3])) + ((s8[4] + s8[5]) + (s8[6] + s8[7]));
        x0[256] = sum;
        __syncthreads();
        sum += y0[256];
        const float inv = sum > 0.f ? 1.f / sum : 0.f;
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c)
#pragma unroll
          for (int j = 0; j < 32; ++j) v[c][j] *= inv;
      } else {
        if (drop_on(p.drop)) {                              // dP = mask * dP_dropped / (1 - p)
          const uint32_t dstep = *p.drop.step;
#pragma unroll
          for (int c = 0; c < Cfg::NC; ++c)
            if (cbeg + c * 32 < p.Lkp) drop_apply32(p.drop, static_cast<unsigned long long>(row_off + cbeg + c * 32), dstep, v[c]);
        }
        // pass 1: dot = sum_j P_j dP_j over the whole row
        const __nv_bfloat16* prow = p.P + row_off;
        float d4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c) {
          const int col0 = cbeg + c * 32;
          if (row_ok && col0 < p.Lk) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (col0 + j >= p.Lkp) break;
              const uint4 w = __ldg(reinterpret_cast<const uint4*>(prow + col0 + j));
              const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = __bfloat1622float2(h2[e]);
                if (col0 + j + 2 * e < p.Lk) d4[e] += f.x * v[c][j + 2 * e];
                if (col0 + j + 2 * e + 1 < p.Lk) d4[e] += f.y * v[c][j + 2 * e + 1];
              }
            }
          }
        }
        float dot = (d4[0] + d4[1]) + (d4[2] + d4[3]);
        *x0 = dot;
        __syncthreads();
        if (leader) next_tile(t, it);
        __syncwarp();
        dot += *y0;
        // pass 2: dS = scale * P * (dP - dot); P is re-read (L1 / L2 hit)
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c) {
          const int col0 = cbeg + c * 32;
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float pv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (row_ok && col0 + j < p.Lkp && col0 < p.Lk) {
              const uint4 w = __ldg(reinterpret_cast<const uint4*>(prow + col0 + j));
              const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
              for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(h2[e]); pv[2 * e] = f.x; pv[2 * e + 1] = f.y; }
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) v[c][j + e] = (col0 + j + e < p.Lk) ? p.scale * pv[e] * (v[c][j + e] - dot) : 0.f;
          }
        }
        __syncthreads();      // keep the two exchange passes of consecutive tiles apart
      }
      // store bf16, pad columns [Lk, Lkp) as zeros (FWD: exp(-inf) = 0 already)
      if (row_ok) {
        __nv_bfloat16* orow = p.out + row_off;
#pragma unroll
        for (int c = 0; c < Cfg::NC; ++c) attn_store_chunk(orow, cbeg + c * 32, v[c], p.Lkp, p.vec32);
        if (!BWD && drop_on(p.drop)) {                      // second output: the dropped weights that multiply V
          const uint32_t dstep = *p.drop.step;
          __nv_bfloat16* drow = p.out2 + row_off;
#pragma unroll
          for (int c = 0; c < Cfg::NC; ++c) {
            if (cbeg + c * 32 >= p.Lkp) continue;
            drop_apply32(p.drop, static_cast<unsigned long long>(row_off + cbeg + c * 32), dstep, v[c]);
            attn_store_chunk(drow, cbeg + c * 32, v[c], p.Lkp, p.vec32);
          }
        }
      }
      if (!BWD) __syncthreads();  // exchange slots are reused by the next tile
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}


// ------------------------------------------------------------------------------------------------------------------
// Long rows (320 < Lk <= 1024 keys: the 3-camera encoder, 902 tokens): the score row no longer fits TMEM + registers at
// once, so the tile is processed in chunks of 256 keys, twice (the 64-deep QK^T is cheap to recompute; the fp32 score
// matrix still never touches HBM -- the unfused path wrote it, re-read it in a softmax kernel and wrote P):
//   pass 1: online max / sum (FWD) or dot = sum_j P_j dP_j (BWD) over the chunks,
//   pass 2: recompute the chunk, normalise -> P bf16 (FWD) / dS = scale * P o (dP - dot) (BWD).
// Work item = (batch, head, group of consecutive 128-query tiles): all keys of that (batch, head) are loaded ONCE per item
// (up to 4 TMA boxes of 256 x 64) and stay resident while the item's query tiles stream through a 2-slot ring.
// Two [128 x 256] fp32 accumulators in TMEM (512 columns): the MMA of chunk i+2 is issued as soon as chunk i sits in
// registers, so the tensor pipe runs under the softmax arithmetic.  Eight warps, all epilogue (two per TMEM lane
// quadrant, half a chunk = 128 columns each); thread 0 doubles as TMA producer and MMA issuer.
__device__ __forceinline__ float fast_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
struct AttnCCfg {
  static constexpr int CH = 256;
  static constexpr int NCH_MAX = 4;
  static constexpr int A_BYTES = 128 * 128;
  static constexpr int B_CH_BYTES = CH * 128;
  static constexpr int PARTS = 4;                       // threads per query row: 16 warps, four per TMEM lane quadrant
  static constexpr int PCOLS = CH / PARTS;              // columns of a chunk per thread (64)
  static constexpr int SMEM = 1024 + 2 * A_BYTES + NCH_MAX * B_CH_BYTES + 2 * PARTS * 128 * 4 + 8 * 8 + 16;
};

template <int BWD>
__global__ void __launch_bounds__(512, 1)
attn_scores_chunked_kernel(const __grid_constant__ CUtensorMap amap, const __grid_constant__ CUtensorMap bmap,
                           const __grid_constant__ AttnParams p) {
  using Cfg = AttnCCfg;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* qbuf = smem;                                                 // [2][128 x 64] query tiles (A operand)
  uint8_t* kbuf = smem + 2 * Cfg::A_BYTES;                              // [nch][256 x 64] keys of this (batch, head)
  float* xch = reinterpret_cast<float*>(kbuf + Cfg::NCH_MAX * Cfg::B_CH_BYTES);   // [2 values][4 parts][128 rows]
  uint64_t* qfull = reinterpret_cast<uint64_t*>(xch + 2 * Cfg::PARTS * 128);
  uint64_t* kfull = qfull + 2;
  uint64_t* tfull = kfull + 1;
  uint64_t* tempty = tfull + 2;                                         // one arrival per warp once its chunk sits in registers
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool leader = threadIdx.x == 0;
  const int pairs = p.heads * p.batch;
  const int items = pairs * p.groups;
  const int nch = p.nch;
  const bool single = BWD && p.O != nullptr;                            // dot = dO . O: no first pass
  const int total = single ? nch : 2 * nch;
  constexpr uint32_t idesc = make_idesc_bf16(128, Cfg::CH, 0, 0);

  pdl_trigger();
  if (leader) {
    prefetch_tmap(&amap);
    prefetch_tmap(&bmap);
    mbar_init(&qfull[0], 1); mbar_init(&qfull[1], 1); mbar_init(kfull, 1);
    mbar_init(&tfull[0], 1); mbar_init(&tfull[1], 1);
    mbar_init(&tempty[0], 4 * Cfg::PARTS); mbar_init(&tempty[1], 4 * Cfg::PARTS);
    fence_mbar_init();
  }
  if (warp == 0) { __syncwarp(); tmem_alloc(tmem_slot, 512); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  const int q = warp & 3, part = warp >> 2;               // TMEM lane quadrant / quarter of a chunk's columns
  const float sl2 = p.scale * 1.4426950408889634f;       // forward works in the log2 domain: exp(x) = ex2(x * log2 e)
  const int r = q * 32 + lane;
  const uint64_t adesc0 = make_smem_desc(smem_u32(qbuf), 16, 1024);
  const uint64_t bdesc0 = make_smem_desc(smem_u32(kbuf), 16, 1024);
  uint32_t nql = 0, nqw = 0, nkw = 0, use0 = 0, use1 = 0;               // Q loads issued / waited, K waits, TMEM buffer uses
  const uint32_t dstep = drop_on(p.drop) ? *p.drop.step : 0u;

  auto mma_step = [&](int slot, int i) {                                // leader: chunk-step i of the current tile
    const int c = i % nch;
    const uint64_t ad0 = adesc0 + static_cast<uint32_t>(slot) * (Cfg::A_BYTES >> 4);
    const uint64_t bd0 = bdesc0 + static_cast<uint32_t>(c) * (Cfg::B_CH_BYTES >> 4);
    const uint32_t td = tmem_base + (i & 1) * Cfg::CH;
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_bf16(td, ad0 + 2 * k, bd0 + 2 * k, idesc, k > 0 ? 1u : 0u);
    umma_commit(&tfull[i & 1]);
  };

  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int pair = item / p.groups, grp = item % p.groups;
    const int hd = pair % p.heads, b = pair / p.heads;
    const int qt0 = grp * p.tiles_per_group, qt1 = min(p.qtiles, qt0 + p.tiles_per_group);
    if (qt0 >= qt1) continue;
    if (leader) {
      // every MMA of the previous item has completed (its last chunk was waited for below): K / Q slots are free
      mbar_expect_tx(kfull, static_cast<uint32_t>(nch) * Cfg::B_CH_BYTES);
      for (int c = 0; c < nch; ++c) tma_load_4d(kbuf + c * Cfg::B_CH_BYTES, &bmap, kfull, 0, c * Cfg::CH, hd, b);   // keys >= Lk: zero fill
      mbar_expect_tx(&qfull[nql & 1], Cfg::A_BYTES);
      tma_load_4d(qbuf + (nql & 1) * Cfg::A_BYTES, &amap, &qfull[nql & 1], 0, qt0 * 128, hd, b);
      ++nql;
    }
    const uint8_t* mrow = (!BWD && p.mask) ? p.mask + static_cast<long long>(b) * p.Lk : nullptr;
    for (int qt = qt0; qt < qt1; ++qt) {
      const int slot = nqw & 1;
      if (leader) {
        if (qt + 1 < qt1) {                                             // next query tile into the other slot
          mbar_expect_tx(&qfull[nql & 1], Cfg::A_BYTES);
          tma_load_4d(qbuf + (nql & 1) * Cfg::A_BYTES, &amap, &qfull[nql & 1], 0, (qt + 1) * 128, hd, b);
          ++nql;
        }
        if (qt == qt0) mbar_wait(kfull, nkw & 1);
        mbar_wait(&qfull[slot], (nqw >> 1) & 1);
        tc_fence_after();
        mma_step(slot, 0);
        mma_step(slot, 1);
      }
      if (qt == qt0) ++nkw;
      ++nqw;
      __syncwarp();
      const int gm = qt * 128 + r;
      const bool row_ok = gm < p.Lq;
      const long long row_off = ((static_cast<long long>(b) * p.heads + hd) * p.Lq + (row_ok ? gm : 0)) * p.Lkp;
      float m_run = -INFINITY, l_run = 0.f, m_tot = 0.f, inv = 0.f, dot = 0.f;
      if (single) {
        // D_i = sum_d dO[i, d] * O[i, d] over the head's 64 dims: 16 per part, met in shared memory
        float part_dot = 0.f;
        if (row_ok) {
          const __nv_bfloat16* orow = p.O + static_cast<long long>(b) * p.o_sz1 + static_cast<long long>(hd) * p.o_sz0 + static_cast<long long>(gm) * p.o_ld + part * 16;
          const __nv_bfloat16* grow = p.dO + static_cast<long long>(b) * p.do_sz1 + static_cast<long long>(hd) * p.do_sz0 + static_cast<long long>(gm) * p.do_ld + part * 16;
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const uint4 wo = __ldg(reinterpret_cast<const uint4*>(orow) + e), wg = __ldg(reinterpret_cast<const uint4*>(grow) + e);
            const __nv_bfloat162* ho = reinterpret_cast<const __nv_bfloat162*>(&wo);
            const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&wg);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 fo = __bfloat1622float2(ho[t]), fg = __bfloat1622float2(hg[t]);
              part_dot += fo.x * fg.x + fo.y * fg.y;
            }
          }
        }
        __syncthreads();                                                // the previous tile's readers of xch are done
        xch[part * 128 + r] = part_dot;
        __syncthreads();
#pragma unroll
        for (int po = 0; po < Cfg::PARTS; ++po) dot += xch[po * 128 + r];
      }
      for (int i = 0; i < total; ++i) {
        const int c = i % nch, pass = single ? 1 : i / nch, buf = i & 1;
        mbar_wait(&tfull[buf], (buf ? use1 : use0) & 1);
        if (buf) ++use1; else ++use0;
        tc_fence_after();
        const uint32_t tacc = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * Cfg::CH + part * Cfg::PCOLS;
        float v[2][32];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          uint32_t u[32];
          tmem_ld_32x32(tacc + k * 32, u);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[k][j] = __uint_as_float(u[j]);
        }
        tc_fence_before();
        // the accumulator buffer is free once every warp holds its part of chunk i in registers: the warps do not wait
        // for each other (no CTA barrier per chunk), only the issuing thread waits for the sixteen arrivals
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[buf]);
        if (leader) {
          mbar_wait(&tempty[buf], ((buf ? use1 : use0) - 1) & 1);
          if (i + 2 < total) { tc_fence_after(); mma_step(slot, i + 2); }
        }
        __syncwarp();
        if (!single && i == nch) __syncthreads();                       // every part's pass-1 partials are in xch
        const int colb = c * Cfg::CH + part * Cfg::PCOLS;
        if (pass == 1 && c == 0 && !single) {                           // the other parts' pass-1 partials (written before the barrier above)
          // (every part of a row combines the four partials in the same order: identical m_tot / inv / dot in all of them)
          if (!BWD) {
            m_tot = -INFINITY;
#pragma unroll
            for (int po = 0; po < Cfg::PARTS; ++po) m_tot = fmaxf(m_tot, xch[po * 128 + r]);
            if (m_tot == -INFINITY) { m_tot = 0.f; inv = 0.f; }
            else {
              float l = 0.f;
#pragma unroll
              for (int po = 0; po < Cfg::PARTS; ++po) l += xch[Cfg::PARTS * 128 + po * 128 + r] * fast_ex2(xch[po * 128 + r] - m_tot);
              inv = l > 0.f ? 1.f / l : 0.f;
            }
          } else {
            dot = 0.f;
#pragma unroll
            for (int po = 0; po < Cfg::PARTS; ++po) dot += xch[po * 128 + r];
          }
        }
        if (!BWD) {
          if (pass == 0) {
            float m8[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) m8[e] = -INFINITY;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const int col0 = colb + k * 32;
              if (!mrow && col0 + 32 <= p.Lk) {
#pragma unroll
                for (int j = 0; j < 32; ++j) { v[k][j] *= sl2; m8[j & 7] = fmaxf(m8[j & 7], v[k][j]); }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  const int col = col0 + j;
                  const bool dead = col >= p.Lk || (mrow && __ldg(mrow + min(col, p.Lk - 1)));
                  v[k][j] = dead ? -INFINITY : v[k][j] * sl2;
                  m8[j & 7] = fmaxf(m8[j & 7], v[k][j]);
                }
              }
            }
            const float cm = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])), fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
            const float mn = fmaxf(m_run, cm);
            if (mn > -INFINITY) {
              float s8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
              for (int k = 0; k < 2; ++k)
#pragma unroll
                for (int j = 0; j < 32; ++j) s8[j & 7] += fast_ex2(v[k][j] - mn);
              l_run = l_run * fast_ex2(m_run - mn) + (((s8[0] + s8[1]) + (s8[2] + s8[3])) + ((s8[4] + s8[5]) + (s8[6] + s8[7])));
              m_run = mn;
            }
            if (c == nch - 1) { xch[part * 128 + r] = m_run; xch[Cfg::PARTS * 128 + part * 128 + r] = l_run; }
          } else {
            const float nm = -m_tot;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const int col0 = colb + k * 32;
              if (!mrow && col0 + 32 <= p.Lk) {                           // interior: no per-element liveness test
#pragma unroll
                for (int j = 0; j < 32; ++j) v[k][j] = fast_ex2(fmaf(v[k][j], sl2, nm)) * inv;
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  const int col = col0 + j;
                  const bool dead = col >= p.Lk || (mrow && __ldg(mrow + min(col, p.Lk - 1)));
                  v[k][j] = dead ? 0.f : fast_ex2(fmaf(v[k][j], sl2, nm)) * inv;
                }
              }
            }
            if (row_ok) {
#pragma unroll
              for (int k = 0; k < 2; ++k) attn_store_chunk(p.out + row_off, colb + k * 32, v[k], p.Lkp, p.vec32);
              if (drop_on(p.drop)) {
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                  if (colb + k * 32 >= p.Lkp) continue;
                  drop_apply32(p.drop, static_cast<unsigned long long>(row_off + colb + k * 32), dstep, v[k]);
                  attn_store_chunk(p.out2 + row_off, colb + k * 32, v[k], p.Lkp, p.vec32);
                }
              }
            }
          }
        } else {
          const __nv_bfloat16* prow = p.P + row_off;
          if (drop_on(p.drop)) {
#pragma unroll
            for (int k = 0; k < 2; ++k)
              if (colb + k * 32 < p.Lkp) drop_apply32(p.drop, static_cast<unsigned long long>(row_off + colb + k * 32), dstep, v[k]);
          }
          float d4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const int col0 = colb + k * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              float pv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              if (row_ok && col0 + j < p.Lkp) {
                const uint4 w = __ldg(reinterpret_cast<const uint4*>(prow + col0 + j));
                const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
                for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(h2[e]); pv[2 * e] = f.x; pv[2 * e + 1] = f.y; }
              }
              if (pass == 0) {
#pragma unroll
                for (int e = 0; e < 8; ++e) if (col0 + j + e < p.Lk) d4[e & 3] += pv[e] * v[k][j + e];
              } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[k][j + e] = (col0 + j + e < p.Lk) ? p.scale * pv[e] * (v[k][j + e] - dot) : 0.f;
              }
            }
          }
          if (pass == 0) {
            dot += (d4[0] + d4[1]) + (d4[2] + d4[3]);
            if (c == nch - 1) { xch[part * 128 + r] = dot; }
          } else if (row_ok) {
#pragma unroll
            for (int k = 0; k < 2; ++k) attn_store_chunk(p.out + row_off, colb + k * 32, v[k], p.Lkp, p.vec32);
          }
        }
      }
    }
    __syncthreads();        // every thread has consumed the item's last chunk before the leader overwrites K / Q
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace act
