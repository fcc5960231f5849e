// K4: perspective writer.  generatePerspective* (src/util.py:46-150, src/numba/util.py:28-76) +
// generatePerspectiveBatch + np.concatenate + float32 cast (src/numba/util_actor.py:33-39, :56-67)
// as one streaming-store kernel.
//
// One warp per environment.  The env's output is one contiguous run of P * 2d^2 floats starting at
// float index offsets[e] * 2d^2 (always even, so 8-byte aligned; 16-byte aligned after peeling a
// 2-float "phantom" head).  The run is produced as a BIT stream first: lane L assembles the 32-bit word
// whose bit t is output float t, by OR-ing the (<= ceil(31/d)+1) d-bit row fields that overlap it.  A row
// field of perspective (m,i,j), plane c, output row a is a rotated row of a 4-entry lookup table kept in
// shared memory (closed forms in oracle/toric_oracle.py:perspective_view):
//   m=0: out[c,a,:] = rotr_d( S[c][(a + i-g) % d]          , (j-g) % d )
//   m=1: out[0,a,:] = rotr_d( V^T [(d - (a + g-j)) % d]    , (i-g) % d )
//        out[1,a,:] = rotr_d( Pl^T[ d-1 - (a + g-j) % d ]  , (i-g) % d )
// where S^T is the bit-transpose (built with d ballots).  With the m=1 tables stored pre-permuted, both
// cases read tab[m][c][a + ra] with ra = (i-g)%d or (g-j)%d, rows duplicated so no modulo is needed,
// and each row stored "doubled" (w | w<<d) so the rotation is one shift.
// The words are then expanded 4 bits -> float4 with warp shuffles so that every store instruction of a
// warp covers 512 contiguous, 16-byte aligned bytes.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace tc {

template <int D>
struct PerspCfg {
  static constexpr bool kWide = (2 * D > 32);
  using TabT = typename std::conditional<kWide, unsigned long long, uint32_t>::type;
  static constexpr int kWarps = (D <= 16) ? 8 : 4;
  static constexpr int kPMax = 2 * D * D;
  static constexpr int kFields = (31 + D - 1) / D + 1;   // max row fields overlapping one 32-bit word
};

__device__ __forceinline__ void st_f4(float* p, float a, float b, float c, float d) {
  __stcs(reinterpret_cast<float4*>(p), make_float4(a, b, c, d));
}
__device__ __forceinline__ void st_f2(float* p, float a, float b) {
  __stcs(reinterpret_cast<float2*>(p), make_float2(a, b));
}

template <int D>
__global__ void __launch_bounds__(PerspCfg<D>::kWarps * 32)
persp_write_kernel(const uint32_t* __restrict__ syn, const int32_t* __restrict__ local_off,
                   const int64_t* __restrict__ block_off, int epb, float* __restrict__ out,
                   uint8_t* __restrict__ pos, int64_t* __restrict__ offsets, int32_t* __restrict__ overflow,
                   int64_t cap, int64_t n) {
  using Cfg = PerspCfg<D>;
  using TabT = typename Cfg::TabT;
  constexpr int G = D / 2;                       // grid_shift (src/Actor_mp.py:59)
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr uint32_t kRowMask = D >= 32 ? 0xffffffffu : ((1u << (D & 31)) - 1u);
  __shared__ uint16_t s_pos[Cfg::kWarps][Cfg::kPMax + 2];
  __shared__ TabT s_tab[Cfg::kWarps][4][2 * D];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t e = (int64_t)blockIdx.x * Cfg::kWarps + warp;
  if (e >= n) return;

  uint32_t v = 0, pl = 0;
  if (lane < D) {
    v = __ldg(syn + e * 2 * D + lane);
    pl = __ldg(syn + e * 2 * D + D + lane);
  }
  uint32_t m0, m1;
  persp_masks<32>(kFull, D, lane, v, pl, m0, m1);
  const int c0 = __popc(m0), c1 = __popc(m1);
  int t0, t1;
  const int b0 = gscan_excl<32>(kFull, c0, lane, t0);
  const int b1 = gscan_excl<32>(kFull, c1, lane, t1) + t0;
  const int P = t0 + t1;
  const int64_t off = block_off[e / epb] + (int64_t)local_off[e];
  if (lane == 0) offsets[e] = off;
  if (P == 0) return;
  if (off + P > cap) {
    if (lane == 0 && overflow) *overflow = 1;
    return;
  }

  // ---- position list (reference order: matrix 0 row-major, then matrix 1 row-major) ----
  {
    uint32_t* pos32 = reinterpret_cast<uint32_t*>(pos);
    const int ia = (lane - G + D) % D;           // (i-g) % d for this lane's row
    uint32_t mm = m0;
    int k = b0;
    while (mm) {
      const int j = __ffs(mm) - 1;
      mm &= mm - 1;
      const int rot = (j - G + D) % D;
      s_pos[warp][k] = (uint16_t)(0 | (ia << 1) | (rot << 6));
      if (pos) pos32[off + k] = 0u | ((uint32_t)lane << 8) | ((uint32_t)j << 16);
      ++k;
    }
    mm = m1;
    k = b1;
    while (mm) {
      const int j = __ffs(mm) - 1;
      mm &= mm - 1;
      const int ra = (G - j + D) % D;
      s_pos[warp][k] = (uint16_t)(1 | (ra << 1) | (ia << 6));
      if (pos) pos32[off + k] = 1u | ((uint32_t)lane << 8) | ((uint32_t)j << 16);
      ++k;
    }
  }

  // ---- lookup tables ----
  {
    uint32_t tv = 0, tp = 0;
#pragma unroll
    for (int b = 0; b < D; ++b) {
      const uint32_t bv = __ballot_sync(kFull, (v >> b) & 1u);
      const uint32_t bp = __ballot_sync(kFull, (pl >> b) & 1u);
      if (lane == b) { tv = bv; tp = bp; }
    }
    if (lane < D) {
      const TabT dv = (TabT)v | ((TabT)v << D), dp = (TabT)pl | ((TabT)pl << D);
      const TabT dtv = (TabT)tv | ((TabT)tv << D), dtp = (TabT)tp | ((TabT)tp << D);
      s_tab[warp][0][lane] = dv;
      s_tab[warp][0][lane + D] = dv;
      s_tab[warp][1][lane] = dp;
      s_tab[warp][1][lane + D] = dp;
      const int kv = (D - lane) % D, kp = D - 1 - lane;
      s_tab[warp][2][kv] = dtv;
      s_tab[warp][2][kv + D] = dtv;
      s_tab[warp][3][kp] = dtp;
      s_tab[warp][3][kp + D] = dtp;
    }
  }
  __syncwarp();

  // ---- bit stream -> float stream ----
  const int64_t f_start = off * (int64_t)(2 * D * D);
  const int ph = (int)(f_start & 3);             // 0 or 2 phantom floats in front (owned by the previous env)
  float* __restrict__ base = out + (f_start - ph);
  const int F = P * 2 * D * D;
  const int L = F + ph;
  const int nW = (L + 31) >> 5;
  const uint16_t* sp = s_pos[warp];
  const TabT* tab = &s_tab[warp][0][0];

  for (int w0 = 0; w0 < nW; w0 += 32) {
    const int W = w0 + lane;
    uint32_t word = 0;
    if (W < nW) {
      const int tbit0 = 32 * W - ph;             // stream bit held by bit 0 of this word
      const int tb = tbit0 < 0 ? 0 : tbit0;
      const int f = tb / D;                      // first row field touching the word
      int p = f / (2 * D);
      int rr = f - p * (2 * D);                  // c * D + a
      int sh = f * D - tbit0;                    // where that field starts inside the word (may be < 0)
      uint32_t ent = sp[p];
#pragma unroll
      for (int it = 0; it < Cfg::kFields; ++it) {
        if (sh < 32 && p < P) {
          const int c = rr >= D ? 1 : 0;
          const int a = rr - c * D;
          const TabT tw = tab[(((ent & 1u) << 1) | c) * (2 * D) + a + ((ent >> 1) & 31u)];
          const uint32_t fld = (uint32_t)(tw >> ((ent >> 6) & 31u)) & kRowMask;
          word |= sh >= 0 ? (fld << sh) : (fld >> (-sh));
        }
        sh += D;
        if (++rr == 2 * D) {
          rr = 0;
          ++p;
          ent = sp[p];
        }
      }
    }
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const uint32_t wv = __shfl_sync(kFull, word, s * 4 + (lane >> 3));
      const int fi = 4 * ((w0 + s * 4) * 8 + lane);   // float index inside the aligned stream
      if (fi < L) {
        const uint32_t nib = wv >> ((lane & 7) * 4);
        const float a = (nib & 1u) ? 1.0f : 0.0f, b = (nib & 2u) ? 1.0f : 0.0f;
        const float c = (nib & 4u) ? 1.0f : 0.0f, d = (nib & 8u) ? 1.0f : 0.0f;
        if (fi >= ph && fi + 4 <= L) st_f4(base + fi, a, b, c, d);
        else if (fi < ph) st_f2(base + fi + 2, c, d);
        else st_f2(base + fi, a, b);
      }
    }
  }
}

}  // namespace tc
