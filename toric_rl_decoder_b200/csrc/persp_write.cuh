// K4: perspective writer.  generatePerspective* (src/util.py:46-150, src/numba/util.py:28-76) +
// generatePerspectiveBatch + np.concatenate + float32 cast (src/numba/util_actor.py:33-39, :56-67)
// as one streaming kernel whose global writes are TMA bulk stores.
//
// One warp per environment.  The env's output is one contiguous run of P * 2d^2 floats starting at
// float index offsets[e] * 2d^2 (always even, so 8-byte aligned; 16-byte aligned after peeling a
// 2-float "phantom" head that belongs to the previous env).  The run is produced as a BIT stream first
// (bit t = output float t), 32 words = 1024 floats = one 4 KB tile at a time:
//   1. lane L assembles word L of the tile (two code paths, below);
//   2. the 32 words are expanded 4 bits -> float4 (warp shuffle + a 16-entry float4 lookup table in
//      shared memory) into a 4 KB staging tile in shared memory, 512 contiguous bytes per STS.128;
//   3. one lane pushes the tile with cp.async.bulk.global.shared::cta (1-D TMA bulk store); tiles are
//      double-buffered per warp, so assembling tile t+1 overlaps the store of tile t.
// Measured on B200 (scripts/micro/write_peak*.cu): warp-wide st.global.v4 streams reach ~6.4 TB/s,
// TMA bulk stores of 4-8 KB tiles ~7.0 TB/s (cudaMemset: 7.27 TB/s) -- hence the staging tile.
//
// Word assembly.  A perspective (m,i,j) shows plane c as a 2-D roll of a base plane (closed forms in
// oracle/toric_oracle.py:perspective_view); with g = d/2, ia = (i-g)%d, rot = (j-g)%d, ra = (g-j)%d:
//   m=0: out[c,a,:] = rotr_d( S[c][(a + ia) % d]           , rot )
//   m=1: out[0,a,:] = rotr_d( V^T [(d - (a + ra)) % d]     , ia  )
//        out[1,a,:] = rotr_d( Pl^T[ d-1 - (a + ra) % d ]   , ia  )
// where S^T is the bit-transpose (built with d ballots).  With the m=1 base planes stored pre-permuted,
// every case is "rows (a + r0) % d of base plane t = 2m+c, each rotated right by r1".
//   * d <= 8 ("plane path"): a whole d x d plane fits one 64-bit word.  All 4*d*d rolled planes of the env
//     are precomputed once with whole-plane shifts (row roll = 64-bit rotate by r0*d, column roll = two
//     masked shifts) into shared memory; a stream word is then the funnel of the 2..5 consecutive planes
//     that overlap it.
//   * d > 8 ("row path"): lane L ORs the (<= ceil(31/d)+1) d-bit row fields that overlap its word, each
//     read from row-doubled tables (w | w<<d, rows duplicated) so roll and rotation need no modulo.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace tc {

#ifndef TC_TILE_WPL
#define TC_TILE_WPL 2
#endif
#ifndef TC_STAGE_BUFS
#define TC_STAGE_BUFS 1
#endif
#ifndef TC_WRITE_WARPS
#define TC_WRITE_WARPS 4
#endif
constexpr int kWordsPerLane = TC_TILE_WPL;     // 32-bit stream words per lane and tile
constexpr int kTileFloats = 1024 * kWordsPerLane;
constexpr int kTileBytes = kTileFloats * 4;
constexpr int kStageBufs = TC_STAGE_BUFS;      // staging tiles per warp (1: wait for the TMA read before refill)

template <int D>
struct PerspCfg {
  static constexpr bool kPlane = (D <= 8);
  static constexpr bool kWide = (2 * D > 32);
  using TabT = typename std::conditional<kWide, unsigned long long, uint32_t>::type;
  static constexpr int kWarps = TC_WRITE_WARPS;
  static constexpr int kPMax = 2 * D * D;
  static constexpr int kDD = D * D;
  static constexpr int kFields = (31 + D - 1) / D + 1;          // row fields overlapping one word
  static constexpr int kPlanesPerWord = 1 + (31 + D * D - 1) / (D * D);
  static constexpr int kTabBytes = kPlane ? 4 * D * D * 8 : 4 * 2 * D * (int)sizeof(TabT);
  static constexpr int kPosBytes = ((kPMax + 2) * 2 + 15) & ~15;
  static constexpr int kAuxBytes = kPlane ? (4 * 16 * 4 + 32 * 8) : 0;   // plane path: doubled base rows + row-rolled planes
  // per-warp shared memory: [stage tiles][tab / planes][pos list][aux]
  static constexpr int kWarpBytes = kStageBufs * kTileBytes + ((kTabBytes + 15) & ~15) + kPosBytes + kAuxBytes;
  static constexpr int kSmemBytes = kWarps * kWarpBytes;
};

__device__ __forceinline__ void st_f2(float* p, float a, float b) {
  __stcs(reinterpret_cast<float2*>(p), make_float2(a, b));
}

__device__ __forceinline__ void bulk_store(float* gdst, const void* ssrc, int bytes) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(sa), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// 4 stream bits -> 4 floats: bit * 0x3f800000 (1.0f), with the bit still at its position in the nibble
__device__ __forceinline__ float4 expand_nibble(uint32_t nib) {
  return make_float4(__uint_as_float((nib & 1u) * 0x3f800000u), __uint_as_float((nib & 2u) * 0x1fc00000u),
                     __uint_as_float((nib & 4u) * 0x0fe00000u), __uint_as_float((nib & 8u) * 0x07f00000u));
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
  constexpr int DD = D * D;
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr uint32_t kRowMask = D >= 32 ? 0xffffffffu : ((1u << (D & 31)) - 1u);
  extern __shared__ __align__(128) unsigned char smem[];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t e = (int64_t)blockIdx.x * Cfg::kWarps + warp;
  if (e >= n) return;
  unsigned char* wb = smem + warp * Cfg::kWarpBytes;
  float* stage = reinterpret_cast<float*>(wb);
  unsigned char* tab_raw = wb + kStageBufs * kTileBytes;
  uint16_t* sp = reinterpret_cast<uint16_t*>(tab_raw + ((Cfg::kTabBytes + 15) & ~15));
  unsigned char* aux = reinterpret_cast<unsigned char*>(sp) + Cfg::kPosBytes;

  uint32_t v = 0, pl = 0;
  if (lane < D) {
    v = __ldg(syn + e * 2 * D + lane);
    pl = __ldg(syn + e * 2 * D + D + lane);
  }
  uint32_t m0, m1;
  persp_masks<32>(kFull, D, lane, v, pl, m0, m1);
  const int P = gsum<32>(kFull, __popc(m0) + __popc(m1));
  const int64_t off = block_off[e / epb] + (int64_t)local_off[e];
  if (lane == 0) offsets[e] = off;
  if (P == 0) return;
  if (off + P > cap) {
    if (lane == 0 && overflow) *overflow = 1;
    return;
  }

  // ---- position list (reference order: matrix 0 row-major, then matrix 1 row-major) ----
  // One candidate qubit q = m*d*d + i*d + j per lane and round; selected candidates are compacted in order
  // with a ballot.  sp[k]: plane path = index of the c=0 rolled plane of perspective k (t=2m, r0, r1);
  // row path = m | r0<<1 | r1<<6.   m=0: (r0, r1) = ((i-g)%d, (j-g)%d);  m=1: ((g-j)%d, (i-g)%d).
  {
    uint32_t* pos32 = reinterpret_cast<uint32_t*>(pos);
    int kbase = 0;
#pragma unroll 1
    for (int q0 = 0; q0 < 2 * DD; q0 += 32) {
      const int q = q0 + lane;
      const int m = q >= DD ? 1 : 0;
      const int rem = q - m * DD;
      const int i = rem / D, j = rem - i * D;
      const uint32_t r0m = __shfl_sync(kFull, m0, i & 31), r1m = __shfl_sync(kFull, m1, i & 31);
      const bool sel = q < 2 * DD && (((m ? r1m : r0m) >> j) & 1u);
      const uint32_t bal = __ballot_sync(kFull, sel);
      if (sel) {
        const int k = kbase + __popc(bal & ((1u << lane) - 1u));
        const int ia = (i - G + D) % D;
        const int r0 = m ? (G - j + D) % D : ia;
        const int r1 = m ? ia : (j - G + D) % D;
        sp[k] = Cfg::kPlane ? (uint16_t)(2 * m * DD + r0 * D + r1) : (uint16_t)(m | (r0 << 1) | (r1 << 6));
        if (pos) pos32[off + k] = (uint32_t)m | ((uint32_t)i << 8) | ((uint32_t)j << 16);
      }
      kbase += __popc(bal);
    }
  }

  // ---- base tables: t = 2m + c; row a of table t (see header) ----
  uint32_t tv = 0, tp = 0;
#pragma unroll
  for (int b = 0; b < D; ++b) {
    const uint32_t bv = __ballot_sync(kFull, (v >> b) & 1u);
    const uint32_t bp = __ballot_sync(kFull, (pl >> b) & 1u);
    if (lane == b) { tv = bv; tp = bp; }
  }
  if constexpr (Cfg::kPlane) {
    unsigned long long* planes = reinterpret_cast<unsigned long long*>(tab_raw);
    uint32_t* rows = reinterpret_cast<uint32_t*>(aux);                         // [4][16]: rows doubled
    unsigned long long* xs = reinterpret_cast<unsigned long long*>(aux + 4 * 16 * 4);   // [4*D] row-rolled planes
    if (lane < D) {
      const int kv = (D - lane) % D, kp = D - 1 - lane;
      rows[0 * 16 + lane] = v;
      rows[0 * 16 + lane + D] = v;
      rows[1 * 16 + lane] = pl;
      rows[1 * 16 + lane + D] = pl;
      rows[2 * 16 + kv] = tv & kRowMask;
      rows[2 * 16 + kv + D] = tv & kRowMask;
      rows[3 * 16 + kp] = tp & kRowMask;
      rows[3 * 16 + kp + D] = tp & kRowMask;
    }
    __syncwarp();
    if (lane < 4 * D) {                            // xs[t*D + r0][a] = T[t][(a + r0) % d]
      const int t = lane / D, r0 = lane - t * D;
      unsigned long long x = 0;
#pragma unroll
      for (int a = 0; a < D; ++a) x |= (unsigned long long)rows[t * 16 + a + r0] << (a * D);
      xs[lane] = x;
    }
    __syncwarp();
    constexpr unsigned long long kMaskDD = DD >= 64 ? ~0ull : ((1ull << (DD & 63)) - 1ull);
    unsigned long long rep = 0;
#pragma unroll
    for (int a = 0; a < D; ++a) rep |= 1ull << (a * D);
#pragma unroll 1
    for (int idx = lane; idx < 4 * DD; idx += 32) {
      const int tr = idx / D;                      // t*D + r0
      const int r1 = idx - tr * D;
      const unsigned long long x = xs[tr];
      // column roll of every row at once: out[b] = w[(b + r1) % d]
      const unsigned long long lo = rep * (unsigned long long)((1u << (D - r1)) - 1u);
      planes[idx] = ((x >> r1) & lo) | ((x << (D - r1)) & (kMaskDD ^ lo));
    }
  } else {
    TabT* tab = reinterpret_cast<TabT*>(tab_raw);
    if (lane < D) {
      const TabT dv = (TabT)v | ((TabT)v << D), dp = (TabT)pl | ((TabT)pl << D);
      const TabT dtv = (TabT)tv | ((TabT)tv << D), dtp = (TabT)tp | ((TabT)tp << D);
      tab[0 * 2 * D + lane] = dv;
      tab[0 * 2 * D + lane + D] = dv;
      tab[1 * 2 * D + lane] = dp;
      tab[1 * 2 * D + lane + D] = dp;
      const int kv = (D - lane) % D, kp = D - 1 - lane;
      tab[2 * 2 * D + kv] = dtv;
      tab[2 * 2 * D + kv + D] = dtv;
      tab[3 * 2 * D + kp] = dtp;
      tab[3 * 2 * D + kp + D] = dtp;
    }
  }
  __syncwarp();

  // ---- bit stream -> float tiles -> TMA bulk stores ----
  const int64_t f_start = off * (int64_t)(2 * DD);
  const int ph = (int)(f_start & 3);             // 0 or 2 phantom floats in front (owned by the previous env)
  float* __restrict__ base = out + (f_start - ph);   // 16-byte aligned
  const int F = P * 2 * DD;
  const int L = F + ph;
  const int nW = (L + 31) >> 5;
  const int nT = (L + kTileFloats - 1) / kTileFloats;
  const int L4 = L & ~3;                         // floats covered by whole float4s

  auto assemble = [&](const int W) -> uint32_t {
    uint32_t word = 0;
    if (W < nW) {
      if constexpr (Cfg::kPlane) {
        const unsigned long long* planes = reinterpret_cast<const unsigned long long*>(tab_raw);
        const int gg = 32 * W - ph + DD;         // shifted by one (empty) plane so that it is never negative
        const int kk = gg / DD;
        const int r = gg - kk * DD;              // first stream bit of the word inside plane kk-1
#pragma unroll
        for (int j = 0; j < Cfg::kPlanesPerWord; ++j) {
          const int k = kk - 1 + j;              // plane k = perspective k>>1, syndrome plane k&1
          const int sh = j * DD - r;
          if (k >= 0 && k < 2 * P && sh < 32) {
            const unsigned long long pv = planes[(int)sp[k >> 1] + (k & 1) * DD];
            word |= (j == 0) ? (uint32_t)(pv >> r) : ((uint32_t)pv << sh);
          }
        }
      } else {
        const TabT* tab = reinterpret_cast<const TabT*>(tab_raw);
        const int tbit0 = 32 * W - ph;           // stream bit held by bit 0 of this word
        const int tb = tbit0 < 0 ? 0 : tbit0;
        const int f = tb / D;                    // first row field touching the word
        int p = f / (2 * D);
        int rr = f - p * (2 * D);                // c * D + a
        int sh = f * D - tbit0;                  // where that field starts inside the word (may be < 0)
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
    }
    return word;
  };

  for (int t = 0; t < nT; ++t) {
    uint32_t word[kWordsPerLane];
#pragma unroll
    for (int u = 0; u < kWordsPerLane; ++u) word[u] = assemble(32 * (kWordsPerLane * t + u) + lane);
    float* st = stage + (t % kStageBufs) * kTileFloats;
    if (t >= kStageBufs) {                       // the TMA engine must have read this buffer (tile t - kStageBufs)
      if (lane == 0) {
        if (kStageBufs == 1) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      }
      __syncwarp();
    }
    float4* st4 = reinterpret_cast<float4*>(st);
#pragma unroll
    for (int u = 0; u < kWordsPerLane; ++u) {
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const uint32_t wv = __shfl_sync(kFull, word[u], s * 4 + (lane >> 3));
        st4[u * 256 + s * 32 + lane] = expand_nibble(wv >> ((lane & 7) * 4));
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      const int f0 = t * kTileFloats;
      const int lo = (t == 0 && ph) ? 4 : 0;
      int hi = L4 - f0;
      if (hi > kTileFloats) hi = kTileFloats;
      if (t == 0 && ph) st_f2(base + 2, st[2], st[3]);                     // head: floats 2,3 of the first float4
      if ((L & 3) && t == nT - 1) st_f2(base + L - 2, st[L - 2 - f0], st[L - 1 - f0]);   // tail half float4
      if (hi > lo) bulk_store(base + f0 + lo, st + lo, (hi - lo) * 4);
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace tc
