// 3x3 / stride-1 convolution (forward and data gradient) with the input halo patch RESIDENT in shared memory.
//
// tc_kernel's MODE_CONV re-fetches the 128-pixel A tile once per filter tap (9x L2->SM traffic, which is what bounds
// it).  Here one TMA box {64 ch, PW = BW+2 pixels, BH+2 rows} per 64-channel block lands the whole halo patch once;
// patch pixel (h', w') sits in smem row h'*PW + w' (128 B, 128B-swizzled).  Accumulator row p <-> output pixel
// (p / PW, p % PW) of the tile, so the A operand of tap (dh, dw) is simply the SAME patch viewed from row
// (dh+1)*PW + (dw+1): a K-major descriptor whose start address is shifted by whole 128-byte rows.
// Columns w >= BW of each patch row produce garbage accumulator rows that the epilogue masks (utilisation BW/PW).
// Weights stream through their own ring, one {BN x 64} block per (channel block, tap).
// Persistent, warp-specialised like tc_kernel: warp 0 TMA, warp 1 MMA issue, warps 2-5 epilogue, 2 TMEM buffers.
#pragma once
#include "tc_kernel.cuh"

namespace act {

template <int BN>
struct Conv3Cfg {
  static constexpr int PATCH_BYTES = 28 * 1024;                   // (BH+2)*PW + 2 rows of 128 B, <= 222 rows
  static constexpr int W_BYTES = BN * 128;
  static constexpr int NP = BN == 64 ? 3 : 2;                     // patch stages
  static constexpr int NW = BN == 64 ? 12 : BN == 128 ? 8 : 4;    // weight stages
  static constexpr int TILE_BYTES = NP * PATCH_BYTES + NW * W_BYTES;
  static constexpr int NBAR = 2 * NP + 2 * NW + 4;
  static constexpr int SMEM = 1024 + TILE_BYTES + NBAR * 8 + 16;
  static constexpr int TMEM_COLS = 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;
};

template <int BN>
__global__ void __launch_bounds__(320, 1)
conv3_kernel(const __grid_constant__ CUtensorMap amap, const __grid_constant__ CUtensorMap bmap,
             const __grid_constant__ TcParams p) {
  using Cfg = Conv3Cfg<BN>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wsm = smem + Cfg::NP * Cfg::PATCH_BYTES;
  uint64_t* pfull = reinterpret_cast<uint64_t*>(smem + Cfg::TILE_BYTES);
  uint64_t* pempty = pfull + Cfg::NP;
  uint64_t* wfull = pempty + Cfg::NP;
  uint64_t* wempty = wfull + Cfg::NW;
  uint64_t* tfull = wempty + Cfg::NW;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.gx * p.gy;

  if (!p.pdl_late) pdl_trigger();              // (see tc_kernel: the late trigger keeps dependents from squatting on SMs)
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&amap);
    prefetch_tmap(&bmap);
    for (int s = 0; s < Cfg::NP; ++s) { mbar_init(&pfull[s], 1); mbar_init(&pempty[s], 1); }
    for (int s = 0; s < Cfg::NW; ++s) { mbar_init(&wfull[s], 1); mbar_init(&wempty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull[b], 1); mbar_init(&tempty[b], 8); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  const uint32_t patch_tx = static_cast<uint32_t>((p.BH + 2) * p.PW) * 128u;

  if (warp == 0) {
    // ============================================================== TMA producer
    if (elect_one()) {
      uint32_t ps = 0, pph = 0, ws = 0, wph = 0;          // ring positions / phases (no div/mod in the loops)
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int n0 = (t % p.gx) * BN;
        int u = t / p.gx;
        const int tw = u % p.tiles_w; u /= p.tiles_w;
        const int th = u % p.tiles_h, tn = u / p.tiles_h;
        for (int cb = 0; cb < p.cin_blocks; ++cb) {
          mbar_wait(&pempty[ps], pph ^ 1);
          mbar_expect_tx(&pfull[ps], patch_tx);
          tma_load_4d(smem + ps * Cfg::PATCH_BYTES, &amap, &pfull[ps], cb * 64, tw * p.BW - 1, th * p.BH - 1, tn);
          if (++ps == Cfg::NP) { ps = 0; pph ^= 1; }
          for (int tap = 0; tap < p.ntaps; ++tap) {
            if ((p.skip_mask >> (tap * p.cin_blocks + cb)) & 1u) continue;     // structurally zero weight block
            mbar_wait(&wempty[ws], wph ^ 1);
            mbar_expect_tx(&wfull[ws], Cfg::W_BYTES);
            tma_load_2d(wsm + ws * Cfg::W_BYTES, &bmap, &wfull[ws], (p.taps[tap].wslot * p.cin_blocks + cb) * 64, n0);
            if (++ws == Cfg::NW) { ws = 0; wph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================================================== MMA issuer
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN, 0, 0);
      const uint32_t sbase = smem_u32(smem), wbase = smem_u32(wsm);
      uint32_t ps = 0, pph = 0, ws = 0, wph = 0, lt = 0;
      const uint64_t adesc0 = make_smem_desc(sbase, 16, 1024), bdesc0 = make_smem_desc(wbase, 16, 1024);
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
        const uint32_t acc = lt & 1;
        mbar_wait(&tempty[acc], ((lt >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        uint32_t first = 1;
        for (int cb = 0; cb < p.cin_blocks; ++cb) {
          mbar_wait(&pfull[ps], pph);
          tc_fence_after();
          const uint64_t pdesc = adesc0 + ps * (Cfg::PATCH_BYTES >> 4);
          for (int tap = 0; tap < p.ntaps; ++tap) {
            if ((p.skip_mask >> (tap * p.cin_blocks + cb)) & 1u) continue;
            mbar_wait(&wfull[ws], wph);
            tc_fence_after();
            const int off = (p.taps[tap].dh + 1) * p.PW + (p.taps[tap].dw + 1);       // patch row of accumulator row 0
            // row-shifted view of the patch: plain descriptor, no base-offset field -- measured on B200 the 128B swizzle
            // is a function of the absolute shared-memory address bits, so TMA's placement and the MMA read agree.
            // Descriptors differ from the stage-0 ones only in the start-address field (16-byte units, never carries out).
            uint64_t ad = pdesc + static_cast<uint32_t>(off) * 8u;
            uint64_t bd = bdesc0 + ws * (Cfg::W_BYTES >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k, ad += 2, bd += 2) {
              umma_bf16(tmem_d, ad, bd, idesc, first ? 0u : 1u);
              first = 0;
            }
            umma_commit(&wempty[ws]);
            if (++ws == Cfg::NW) { ws = 0; wph ^= 1; }
          }
          umma_commit(&pempty[ps]);
          if (++ps == Cfg::NP) { ps = 0; pph ^= 1; }
        }
        umma_commit(&tfull[acc]);
      }
      if (p.pdl_late) pdl_trigger();
    }
  } else {
    // ============================================================== epilogue warps (2..9): two per TMEM lane quadrant
    const int q = warp & 3;
    const int cbeg = ((warp - 2) >> 2) * (BN / 2), cend = cbeg + BN / 2;
    const int r = q * 32 + lane;                       // accumulator row = patch-linear output position
    const int ph = r / p.PW, pw = r % p.PW;
    uint32_t lt = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++lt) {
      const int n0 = (t % p.gx) * BN;
      int u = t / p.gx;
      const int tw = u % p.tiles_w; u /= p.tiles_w;
      const int th = u % p.tiles_h, tn = u / p.tiles_h;
      const uint32_t acc = lt & 1;
      const int gw = tw * p.BW + pw, gh = th * p.BH + ph;
      const bool row_ok = ph < p.BH && pw < p.BW && gw < p.vW && gh < p.vH && tn < p.NIMG;
      const long long o_off = pix_off(p.og, tn, gh, gw), r_off = pix_off(p.rg, tn, gh, gw);
      mbar_wait(&tfull[acc], (lt >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_acc = tmem_base + acc * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c0 = cbeg; c0 < cend; c0 += 32) {
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32(tmem_acc + c0, v);
        tmem_ld_wait();
        if (c0 + 32 >= cend) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[acc]);
        }
        if (!row_ok || (p.debug & 1)) continue;
        epi_store_chunk(p, v, n0 + c0, o_off, r_off);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

}  // namespace act
