// Train-mode dropout of the transformer blocks (transformer.py:194,197,202-203,219-222,254-266,284-293; p = 0.1 at
// detr/main.py:45), counter-based so that no mask is ever stored: the keep decision of element `e` of dropout site
// `site` at training step `step` is a pure function Philox4x32-10(key = seed, counter = (e / 8, site, step)), recomputed
// wherever it is needed (forward epilogue, backward).  One Philox call yields 128 bits = eight 16-bit lanes, one per
// element: keep iff lane < thresh, thresh = round((1 - p) * 65536).  The step counter lives in device memory so that a
// captured CUDA graph draws fresh masks on every replay.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace act {

struct DropDesc {
  uint32_t seed_lo = 0, seed_hi = 0;
  uint32_t site = 0;
  uint32_t thresh = 0x10000u;            // >= 0x10000: keep everything (dropout off)
  float scale = 1.f;                     // 1 / (1 - p)
  const uint32_t* step = nullptr;        // device counter, bumped once per training forward
};

__host__ __device__ __forceinline__ bool drop_on(const DropDesc& d) { return d.thresh < 0x10000u; }

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return c;
}

// keep bits of elements [8 g, 8 g + 8) of the site's stream: bit i set <=> element 8 g + i is kept
__device__ __forceinline__ uint32_t drop_keep8(const DropDesc& d, unsigned long long g, uint32_t step) {
  const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), d.site, step), d.seed_lo, d.seed_hi);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m |= ((w[i] & 0xffffu) < d.thresh ? 1u : 0u) << (2 * i);
    m |= ((w[i] >> 16) < d.thresh ? 1u : 0u) << (2 * i + 1);
  }
  return m;
}

// In-place dropout of 32 consecutive elements starting at element index e0 (a multiple of 8).
__device__ __forceinline__ void drop_apply32(const DropDesc& d, unsigned long long e0, uint32_t step, float (&f)[32]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t m = drop_keep8(d, (e0 >> 3) + q, step);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[q * 8 + i] = ((m >> i) & 1u) ? f[q * 8 + i] * d.scale : 0.f;
  }
}

}  // namespace act
