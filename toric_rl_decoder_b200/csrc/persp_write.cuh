// K4: perspective writer.  generatePerspective* (src/util.py:46-150, src/numba/util.py:28-76) +
// generatePerspectiveBatch + np.concatenate + float32 cast (src/numba/util_actor.py:33-39, :56-67)
// as one streaming kernel whose global writes are TMA bulk stores.
//
// One env at a time per warp; the kernel body is a grid-stride loop over envs, so it runs either with one
// env per warp (default: measured fastest -- the hardware CTA scheduler staggers the setup of new envs against
// the streaming of the others; 6.73 TB/s) or with persistent warps (TC_WRITE_WAVES > 0 in the launcher;
// 5.4-6.3 TB/s: the warps of an SM fall into lock step behind the TMA queue).  The env's output is one contiguous
// run of P * 2d^2 floats starting at float index offsets[e] * 2d^2 (always even, so 8-byte aligned;
// 16-byte aligned after peeling a 2-float "phantom" head that belongs to the previous env).  The run is
// produced as a BIT stream first (bit t = output float t), one tile of 32*kWordsPerLane words at a time:
//   1. lane L assembles words L, L+32, .. of the tile (two code paths, below);
//   2. the words are expanded 4 bits -> float4 (warp shuffle + 8 integer ops) into a staging tile in
//      shared memory, 512 contiguous bytes per STS.128;
//   3. one lane pushes the tile with cp.async.bulk.global.shared::cta (1-D TMA bulk store).  Staging tiles
//      are a per-warp ring (kStageBufs) that keeps rotating across envs; before a tile is refilled the warp
//      waits (cp.async.bulk.wait_group.read) until the TMA engine has read it.
// Measured on B200 (scripts/micro/write_peak*.cu): warp-wide st.global.v4 streams reach ~6.4 TB/s,
// TMA bulk stores of 4-8 KB tiles ~7.0 TB/s (cudaMemset: 7.27 TB/s) -- hence the staging tile.  Tile size,
// ring depth and warps per CTA were chosen by A/B timing (scripts/micro/build_variants.py, run_variants.sh):
// 2 warps per CTA, 2 x 4 KB tiles per warp (10.3 KB of shared memory per warp at d=7, 22 warps per SM).
//
// Word assembly.  A perspective (m,i,j) shows plane c as a 2-D roll of a base plane (closed forms in
// oracle/toric_oracle.py:perspective_view); with g = d/2, ia = (i-g)%d, rot = (j-g)%d, ra = (g-j)%d:
//   m=0: out[c,a,:] = rotr_d( S[c][(a + ia) % d]           , rot )
//   m=1: out[0,a,:] = rotr_d( V^T [(d - (a + ra)) % d]     , ia  )
//        out[1,a,:] = rotr_d( Pl^T[ d-1 - (a + ra) % d ]   , ia  )
// where S^T is the bit-transpose (built with d ballots).  With the m=1 base planes stored pre-permuted,
// every case is "rows (a + r0) % d of base plane t = 2m+c, each rotated right by r1".
//   * d <= 8 ("plane path"): a whole d x d plane fits one 64-bit word.  All 4*d*d rolled planes of the env
//     are precomputed once (lane t*d+r0 builds the row-rolled plane, then d column rolls with compile-time
//     masks and shifts) into shared memory; a stream word is then the funnel of the 2..5 consecutive planes
//     that overlap it.  The env-independent decode of a candidate qubit q -> (plane index, position word)
//     is a per-CTA table built once.
//   * d > 8 ("row path"): lane L ORs the (<= ceil(31/d)+1) d-bit row fields that overlap its word, each
//     read from row-doubled tables (w | w<<d, rows duplicated) so roll and rotation need no modulo.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace tc {

#ifndef TC_TILE_WPL
#define TC_TILE_WPL 1
#endif
#ifndef TC_STAGE_BUFS
#define TC_STAGE_BUFS 2
#endif
#ifndef TC_WRITE_WARPS
#define TC_WRITE_WARPS 2
#endif
constexpr int kTileBytes = 4096 * TC_TILE_WPL; // staging tile: 32 lanes x TC_TILE_WPL stream words of float32 output
constexpr int kStageBufsDefault = TC_STAGE_BUFS;   // staging tiles per warp (1: wait for the TMA read before refill)

// Output element types.  The stream is 0/1 valued, so every type is exact; float32 is the reference's
// (`.type('torch.Tensor')`, src/numba/util_actor.py:39), the narrower ones feed a bf16/fp16-autocast or
// integer-input network without a cast kernel.  OutT = float | uint16_t (bf16 / fp16, the bit pattern of 1.0
// is a kernel argument) | uint8_t.  kE = elements per 16-byte chunk = stream bits per STS.128.
template <typename OutT>
struct OutCfg {
  static constexpr int kB = (int)sizeof(OutT);
  static constexpr int kE = 16 / kB;
  static constexpr int kWordsPerLane = TC_TILE_WPL * 4 / kB;   // 32-bit stream words per lane and tile
  static constexpr int kTileElems = 1024 * kWordsPerLane;      // = kTileBytes / kB
  using PairT = typename std::conditional<kB == 4, float2,
                                          typename std::conditional<kB == 2, uint32_t, uint16_t>::type>::type;
};

template <int D>
struct PerspCfg {
  static constexpr bool kPlane = (D <= 8);
  static constexpr bool kWide = (2 * D > 32);
  using TabT = typename std::conditional<kWide, unsigned long long, uint32_t>::type;
  static constexpr int kWarps = TC_WRITE_WARPS;
#ifndef TC_COUNT_THREADS
#define TC_COUNT_THREADS 256
#endif
  static constexpr int kEpb = TC_COUNT_THREADS / (D <= 8 ? 8 : (D <= 16 ? 16 : 32));   // envs per CTA of the count kernels
  static constexpr int kPMax = 2 * D * D;
  static constexpr int kDD = D * D;
  static constexpr int kFields = (31 + D - 1) / D + 1;          // row fields overlapping one word
  static constexpr int kPlanesPerWord = 1 + (31 + D * D - 1) / (D * D);
  // Small lattices: an env's whole output is less than two 4 KB tiles (d=3: ~0.9 KB, d=4: ~3 KB, d=5: ~8 KB), so one
  // warp per env spends its time on per-env setup and on expanding mostly empty tiles.  persp_write_group_kernel
  // gives a warp kGroup CONSECUTIVE envs instead: their outputs are adjacent in memory, so they are written as ONE bit
  // stream (one head / tail, full tiles in between) from kGroup plane tables.
#ifndef TC_GROUP_D3
#define TC_GROUP_D3 8
#endif
#ifndef TC_GROUP_D4
#define TC_GROUP_D4 4
#endif
#ifndef TC_GROUP_D5
#define TC_GROUP_D5 4
#endif
  static constexpr int kGroup = D == 3 ? TC_GROUP_D3 : (D == 4 ? TC_GROUP_D4 : (D == 5 ? TC_GROUP_D5 : 1));
  // load phase: lane = slot * kLanesPerEnv + row; 4 lanes per env where the rows fit (d <= 4) and the group is large
  static constexpr int kLanesPerEnv = (D <= 4 && kGroup >= 8) ? 4 : (kGroup >= 4 ? 8 : (kGroup == 2 ? 16 : 32));
  static constexpr int kLoadRounds = kGroup * kLanesPerEnv > 32 ? kGroup * kLanesPerEnv / 32 : 1;
  static constexpr int kTabBytes = kPlane ? kGroup * 4 * D * D * 8 : 4 * 2 * D * (int)sizeof(TabT);
  static constexpr int kPosBytes = ((kGroup * kPMax + 2) * 2 + 15) & ~15;
  static constexpr int kAuxBytes = kPlane ? 4 * 16 * 4 : 0;      // plane path: the 4 x d base rows, doubled
  static constexpr int kCandBytes = kPlane ? ((2 * D * D * 8 + 127) & ~127) : 0;   // per-CTA candidate table
  // per-warp shared memory: [stage tiles][tab / planes][pos list][aux]
  // d=3: the writer is bound by per-env set-up latency, not by the store stream -- one staging tile per warp buys 25
  // instead of 16 resident warps per SM (+19 %, profiles/r02_group_writer_sizes.txt); everywhere else two tiles win
  static constexpr int kStageBufs = D == 3 ? 1 : kStageBufsDefault;
  static constexpr int kWarpBytes = kStageBufs * kTileBytes + ((kTabBytes + 15) & ~15) + kPosBytes + kAuxBytes;
  static constexpr int kSmemBytes = kCandBytes + kWarps * kWarpBytes;
  // the group writer takes block_off / local_off of its first env only: a group must not straddle two count-CTAs
  static_assert(kEpb % kGroup == 0, "TC_GROUP_D*: the env group of a writer warp must divide the count kernels' envs per CTA");
};

// TC_TMA_L2_HINT: 0 = none (default), 1 = evict_first, 2 = evict_last, 3 = no_allocate-like evict_first at fraction 0
// -- an L2 eviction-priority hint on the bulk stores (the output is never re-read by this kernel); A/B in
// profiles/r02_ab_tma_l2_hint.txt
#ifndef TC_TMA_L2_HINT
#define TC_TMA_L2_HINT 0
#endif
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, int bytes) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(ssrc);
#if TC_TMA_L2_HINT == 0
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(sa), "r"(bytes)
               : "memory");
#else
  unsigned long long pol;
#if TC_TMA_L2_HINT == 1
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
#elif TC_TMA_L2_HINT == 2
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
#else
  asm volatile("createpolicy.fractional.L2::evict_unchanged.b64 %0, 1.0;" : "=l"(pol));
#endif
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst), "r"(sa),
               "r"(bytes), "l"(pol)
               : "memory");
#endif
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// One lane of the (converged) warp, the same one on every call: the bulk async-groups it commits are per-thread
// state, and ptxas knows a single thread is active behind elect.sync, so the uniform-register operands of UBLKCP
// need no per-lane waterfall loop (with `if (lane == 0)` it emits one).
__device__ __forceinline__ bool elect_one() {
#ifdef TC_NO_ELECT
  return (threadIdx.x & 31) == 0;
#else
  uint32_t is_leader;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(is_leader));
  return is_leader != 0;
#endif
}

// kE stream bits -> one 16-byte chunk of output elements
template <typename OutT>
__device__ __forceinline__ uint4 expand_chunk(uint32_t bits, uint32_t one16) {
  if constexpr (sizeof(OutT) == 4) {
    // 4 bits -> 4 floats: bit * 0x3f800000 (1.0f), with the bit still at its position in the nibble
    return make_uint4((bits & 1u) * 0x3f800000u, (bits & 2u) * 0x1fc00000u, (bits & 4u) * 0x0fe00000u,
                      (bits & 8u) * 0x07f00000u);
  } else if constexpr (sizeof(OutT) == 2) {
    // 8 bits -> 8 halves, two per 32-bit word; one16 = bit pattern of 1.0 (0x3f80 bf16, 0x3c00 fp16).
    // A nibble times 0x00204081 lands bit k at bit 8k (the partial products do not overlap); a byte permute
    // then pairs bytes (2k, 2k+1) as the two 0/1 halves of word k, and one multiply scales both halves.
    const uint32_t lo = ((bits & 15u) * 0x00204081u) & 0x01010101u;
    const uint32_t hi = (((bits >> 4) & 15u) * 0x00204081u) & 0x01010101u;
    return make_uint4(__byte_perm(lo, 0u, 0x4140) * one16, __byte_perm(lo, 0u, 0x4342) * one16,
                      __byte_perm(hi, 0u, 0x4140) * one16, __byte_perm(hi, 0u, 0x4342) * one16);
  } else {
    // 16 bits -> 16 bytes: a nibble times 0x00204081 lands bit k at bit 8k (the four partial products do not overlap)
    return make_uint4(((bits & 15u) * 0x00204081u) & 0x01010101u, (((bits >> 4) & 15u) * 0x00204081u) & 0x01010101u,
                      (((bits >> 8) & 15u) * 0x00204081u) & 0x01010101u,
                      (((bits >> 12) & 15u) * 0x00204081u) & 0x01010101u);
  }
}

// Per-warp shared-memory layout of the writer: [stage tiles][tab / planes][pos list][aux]; `cand` is per CTA.
template <int D, typename OutT>
struct WriterCtx {
  OutT* stage;
  unsigned char* tab_raw;
  uint16_t* sp;
  unsigned char* aux;
  const uint2* cand;
  int lane;
  __device__ __forceinline__ WriterCtx(unsigned char* smem, int warp, int lane_) : lane(lane_) {
    using Cfg = PerspCfg<D>;
    unsigned char* wb = smem + Cfg::kCandBytes + warp * Cfg::kWarpBytes;
    stage = reinterpret_cast<OutT*>(wb);
    tab_raw = wb + Cfg::kStageBufs * kTileBytes;
    sp = reinterpret_cast<uint16_t*>(tab_raw + ((Cfg::kTabBytes + 15) & ~15));
    aux = reinterpret_cast<unsigned char*>(sp) + Cfg::kPosBytes;
    cand = reinterpret_cast<const uint2*>(smem);
  }
};

// per-CTA candidate table (plane path): q = m*d*d + i*d + j -> {plane index | i<<16, position word}
// plane index = 2m*d*d + r0*d + r1 with (r0,r1) = ((i-g)%d, (j-g)%d) for m=0 and ((g-j)%d, (i-g)%d) for m=1.
// Ends with a CTA barrier when the table exists.
template <int D>
__device__ __forceinline__ void build_cand_table(unsigned char* smem, int nthreads) {
  using Cfg = PerspCfg<D>;
  constexpr int G = D / 2, DD = D * D;
  if constexpr (Cfg::kPlane) {
    uint2* cand = reinterpret_cast<uint2*>(smem);
    for (int q = threadIdx.x; q < 2 * DD; q += nthreads) {
      const int m = q >= DD ? 1 : 0;
      const int rem = q - m * DD;
      const int i = rem / D, j = rem - i * D;
      const int ia = (i - G + D) % D;
      const int r0 = m ? (G - j + D) % D : ia;
      const int r1 = m ? ia : (j - G + D) % D;
      cand[q] = make_uint2((uint32_t)(2 * m * DD + r0 * D + r1) | ((uint32_t)i << 16) | ((uint32_t)j << 24) |
                               ((uint32_t)m << 31),
                           (uint32_t)m | ((uint32_t)i << 8) | ((uint32_t)j << 16));
    }
    __syncthreads();
  }
}

// Position list of one env (reference order: matrix 0 row-major, then matrix 1 row-major), appended to the warp's
// list at index kbase0; returns the list length afterwards.  slot: which plane table of the warp the env uses (plane
// path with several envs per warp; 0 otherwise).  pos32 (nullable): the global (m,i,j,0) array, off: index of the
// warp's FIRST perspective in it.
template <int D, typename OutT>
__device__ __forceinline__ int env_positions(const WriterCtx<D, OutT>& cx, uint32_t m0, uint32_t m1, int kbase0,
                                             int slot, uint32_t* __restrict__ pos32, int64_t off) {
  using Cfg = PerspCfg<D>;
  constexpr int G = D / 2;                       // grid_shift (src/Actor_mp.py:59)
  constexpr int DD = D * D;
  constexpr uint32_t kFull = 0xffffffffu;
  const int lane = cx.lane;
  uint16_t* const sp = cx.sp;
  const uint2* const cand = cx.cand;
  uint8_t* const pos_w = reinterpret_cast<uint8_t*>(pos32);
  int kend;
  // ---- position list (reference order: matrix 0 row-major, then matrix 1 row-major) ----
  // One candidate qubit per lane and round; selected candidates are compacted in order with a ballot.
  // sp[k]: plane path = index of the c=0 rolled plane of perspective k; row path = m | r0<<1 | r1<<6.
  {
    int kbase = kbase0;
#pragma unroll 1
    for (int q0 = 0; q0 < 2 * DD; q0 += 32) {
      const int q = q0 + lane;
      if constexpr (Cfg::kPlane) {
        const uint2 cd = cand[q < 2 * DD ? q : 0];
        const uint32_t r0m = __shfl_sync(kFull, m0, (cd.x >> 16) & 31u);
        const uint32_t r1m = __shfl_sync(kFull, m1, (cd.x >> 16) & 31u);
        const bool sel = q < 2 * DD && ((((cd.x >> 31) ? r1m : r0m) >> ((cd.x >> 24) & 31u)) & 1u);
        const uint32_t bal = __ballot_sync(kFull, sel);
        if (sel) {
          const int k = kbase + __popc(bal & ((1u << lane) - 1u));
          sp[k] = (uint16_t)(cd.x + (uint32_t)(slot * 4 * DD));
          if (pos_w) pos32[off + k] = cd.y;
        }
        kbase += __popc(bal);
      } else {
        const int m = q >= DD ? 1 : 0;
        const int rem = q - m * DD;
        const int i = rem / D, j = rem - i * D;
        const uint32_t r0m = __shfl_sync(kFull, m0, i & 31), r1m = __shfl_sync(kFull, m1, i & 31);
        const bool sel = q < 2 * DD && (((m ? r1m : r0m) >> j) & 1u);
        const uint32_t bal = __ballot_sync(kFull, sel);
        if (sel) {
          const int k = kbase + __popc(bal & ((1u << lane) - 1u));
          int ia = i - G; if (ia < 0) ia += D;
          int rj = m ? G - j : j - G; if (rj < 0) rj += D;
          const int r0 = m ? rj : ia, r1 = m ? ia : rj;
          sp[k] = (uint16_t)(m | (r0 << 1) | (r1 << 6));
          if (pos_w) pos32[off + k] = (uint32_t)m | ((uint32_t)i << 8) | ((uint32_t)j << 16);
        }
        kbase += __popc(bal);
      }
    }
    kend = kbase;
  }
  return kend;
}

// Base tables of one env: t = 2m + c; row a of table t (see header).  Plane path: the 4*d*d rolled planes of the env
// go to plane table `slot` of the warp.
template <int D, typename OutT>
__device__ __forceinline__ void env_tables(const WriterCtx<D, OutT>& cx, uint32_t v, uint32_t pl, int slot) {
  using Cfg = PerspCfg<D>;
  using TabT = typename Cfg::TabT;
  constexpr int DD = D * D;
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr uint32_t kRowMask = D >= 32 ? 0xffffffffu : ((1u << (D & 31)) - 1u);
  const int lane = cx.lane;
  unsigned char* const tab_raw = cx.tab_raw;
  unsigned char* const aux = cx.aux;
  // ---- base tables: t = 2m + c; row a of table t (see header) ----
  uint32_t tv = 0, tp = 0;
#pragma unroll
  for (int b = 0; b < D; ++b) {
    const uint32_t bv = __ballot_sync(kFull, (v >> b) & 1u);
    const uint32_t bp = __ballot_sync(kFull, (pl >> b) & 1u);
    if (lane == b) { tv = bv; tp = bp; }
  }
  if constexpr (Cfg::kPlane) {
    unsigned long long* planes = reinterpret_cast<unsigned long long*>(tab_raw) + slot * 4 * DD;
    uint32_t* rows = reinterpret_cast<uint32_t*>(aux);                         // [4][16]: rows doubled
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
    if (lane < 4 * D) {
      // lane = t*d + r0: row-rolled plane x[a] = T[t][(a + r0) % d], then the d column rolls
      const int t = lane / D, r0 = lane - t * D;
      unsigned long long x = 0;
#pragma unroll
      for (int a = 0; a < D; ++a) x |= (unsigned long long)rows[t * 16 + a + r0] << (a * D);
      constexpr unsigned long long kMaskDD = DD >= 64 ? ~0ull : ((1ull << (DD & 63)) - 1ull);
      unsigned long long rep = 0;
#pragma unroll
      for (int a = 0; a < D; ++a) rep |= 1ull << (a * D);
#pragma unroll
      for (int r1 = 0; r1 < D; ++r1) {           // out[b] = w[(b + r1) % d] for every row at once
        const unsigned long long lo = rep * (unsigned long long)((1u << (D - r1)) - 1u);
        planes[lane * D + r1] = ((x >> r1) & lo) | ((x << (D - r1)) & (kMaskDD ^ lo));
      }
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
}

// The warp's perspective list (P entries, positions and tables set up) as one bit stream -> staging tiles -> TMA bulk
// stores, starting at perspective index `off` of `out`; (slice, split): this warp's share of the tiles.
template <int D, typename OutT>
__device__ __forceinline__ void stream_tiles(const WriterCtx<D, OutT>& cx, int P, int64_t off, int slice, int split,
                                             uint32_t& ring, OutT* __restrict__ out, uint32_t one16) {
  using Cfg = PerspCfg<D>;
  using OC = OutCfg<OutT>;
  constexpr int kE = OC::kE;
  constexpr int kWordsPerLane = OC::kWordsPerLane;
  constexpr int kTileElems = OC::kTileElems;
  using TabT = typename Cfg::TabT;
  constexpr int DD = D * D;
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr uint32_t kRowMask = D >= 32 ? 0xffffffffu : ((1u << (D & 31)) - 1u);
  const int lane = cx.lane;
  OutT* const stage = cx.stage;
  unsigned char* const tab_raw = cx.tab_raw;
  uint16_t* const sp = cx.sp;
  // ---- bit stream -> output tiles -> TMA bulk stores ----
  const int64_t f_start = off * (int64_t)(2 * DD);
  // phantom elements in front (owned by the previous env; even): the stream starts at the kAlignBytes boundary
  // below the env's first element, so every tile boundary after the first is aligned to it in global memory.
  // Whole 128-byte lines per TMA store are worth +0.6 % (d=7) to +1.1 % (d=9) where the writer is HBM-bound
  // (float32, d >= 6); where it is instruction-bound (small d, narrow elements) the up-to-31 extra phantom
  // elements cost 1-2 %, so those keep the minimal 16-byte alignment (profiles/r01_writer_variants.txt).
#ifdef TC_ALIGN_BYTES
  constexpr int kAlignBytes = TC_ALIGN_BYTES;
#else
  constexpr int kAlignBytes = (OC::kB == 4 && D >= 6) ? 128 : 16;
#endif
  constexpr int kAlignElems = kAlignBytes / OC::kB;
  static_assert(kAlignElems >= kE, "alignment below one 16-byte chunk");
  constexpr int kShiftPlanes = (kAlignElems + DD - 1) / DD;   // empty planes in front so that gg below is >= 0
  const int ph = (int)(f_start & (kAlignElems - 1));
  const int phc = ph & ~(kE - 1);              // start of the 16-byte chunk holding the first own element
  const int phr = ph - phc;                    // own elements start phr into that chunk (partial if != 0)
  OutT* __restrict__ base = out + (f_start - ph);
  const int F = P * 2 * DD;                    // >= 18 > kE, so head and tail chunks are never the same chunk
  const int L = F + ph;
  const int nW = (L + 31) >> 5;
  const int nT = (L + kTileElems - 1) / kTileElems;
  const int L16 = L & ~(kE - 1);               // elements covered by whole 16-byte chunks

  auto assemble = [&](const int W) -> uint32_t {
    uint32_t word = 0;
    if (W < nW) {
      if constexpr (Cfg::kPlane) {
        const unsigned long long* planes = reinterpret_cast<const unsigned long long*>(tab_raw);
        const int gg = 32 * W - ph + kShiftPlanes * DD;   // shifted by empty planes: never negative
        const int kk = gg / DD;
        const int r = gg - kk * DD;            // first stream bit of the word inside plane kk - kShiftPlanes
#pragma unroll
        for (int j = 0; j < Cfg::kPlanesPerWord; ++j) {
          const int k = kk - kShiftPlanes + j;  // plane k = perspective k>>1, syndrome plane k&1
          const int sh = j * DD - r;
          if (k >= 0 && k < 2 * P && sh < 32) {
            const unsigned long long pv = planes[(int)sp[k >> 1] + (k & 1) * DD];
            word |= (j == 0) ? (uint32_t)(pv >> r) : ((uint32_t)pv << sh);
          }
        }
      } else {
        const TabT* tab = reinterpret_cast<const TabT*>(tab_raw);
        const int tbit0 = 32 * W - ph;         // stream bit held by bit 0 of this word
        const int tb = tbit0 < 0 ? 0 : tbit0;
        const int f = tb / D;                  // first row field touching the word
        int p = f / (2 * D);
        int rr = f - p * (2 * D);              // c * D + a
        int sh = f * D - tbit0;                // where that field starts inside the word (may be < 0)
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

  // contiguous tile range per slice (measured: dealing the tiles round-robin to the slices is 15 % SLOWER than
  // one warp per env at d=25, contiguous ranges are 10 % faster -- profiles/r01_writer_split_d25.txt)
  const int t_begin = (int)((int64_t)nT * slice / split);
  const int t_end = (int)((int64_t)nT * (slice + 1) / split);
  for (int t = t_begin; t < t_end; ++t, ++ring) {
    uint32_t word[kWordsPerLane];
#pragma unroll
    for (int u = 0; u < kWordsPerLane; ++u) word[u] = assemble(32 * (kWordsPerLane * t + u) + lane);
    constexpr int kStageBufs = Cfg::kStageBufs;
    OutT* st = stage + (ring % kStageBufs) * kTileElems;
    if (ring >= kStageBufs) {                  // the TMA engine must have read the tile pushed kStageBufs ago
      if (elect_one()) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kStageBufs - 1) : "memory");
      __syncwarp();
    }
    uint4* st4 = reinterpret_cast<uint4*>(st);
    constexpr int kSub = 32 / kE;              // 16-byte chunks per stream word (= lanes sharing a word)
#pragma unroll
    for (int u = 0; u < kWordsPerLane; ++u) {
#pragma unroll
      for (int s = 0; s < kSub; ++s) {         // 32 lanes x 16 bytes = 512 contiguous bytes per STS.128
        const uint32_t wv = __shfl_sync(kFull, word[u], s * kE + lane / kSub);
        st4[(u * kSub + s) * 32 + lane] = expand_chunk<OutT>(wv >> ((lane & (kSub - 1)) * kE), one16);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const int f0 = t * kTileElems;
    // partial 16-byte chunks (shared with the neighbouring envs) go out as element pairs: head on lanes
    // 0..kE/2-1, tail on lanes 8..8+kE/2-1
    using PairT = typename OC::PairT;
    if (t == 0 && phr && lane < kE / 2 && 2 * lane >= phr)
      __stcs(reinterpret_cast<PairT*>(base + phc) + lane, reinterpret_cast<const PairT*>(st + phc)[lane]);
    if (t == nT - 1 && lane >= 8 && 2 * (lane - 8) < L - L16)
      __stcs(reinterpret_cast<PairT*>(base + L16) + (lane - 8),
             reinterpret_cast<const PairT*>(st + (L16 - f0))[lane - 8]);
    if (elect_one()) {
      const int lo = t == 0 ? (phr ? phc + kE : phc) : 0;
      int hi = L16 - f0;
      if (hi > kTileElems) hi = kTileElems;
      if (hi > lo) bulk_store(base + f0 + lo, st + lo, (hi - lo) * OC::kB);
      else asm volatile("cp.async.bulk.commit_group;" ::: "memory");     // keep one group per tile
    }
  }
  __syncwarp();
}

// One env (or one slice of its tiles) by one warp: position list, base tables, stream.  v / pl: the env's syndrome rows
// (lane r holds row r), m0 / m1 its selection masks, P > 0 its perspective count, off its first perspective's index in
// `out`.  pos_w: where positions go (nullptr for slices other than 0).  `ring` counts the staging tiles this warp has
// pushed so far (keeps rotating across envs).
template <int D, typename OutT>
__device__ __forceinline__ void write_env(const WriterCtx<D, OutT>& cx, uint32_t v, uint32_t pl, uint32_t m0,
                                          uint32_t m1, int P, int64_t off, int slice, int split, uint32_t& ring,
                                          OutT* __restrict__ out, uint8_t* __restrict__ pos_w, uint32_t one16) {
  env_positions<D, OutT>(cx, m0, m1, 0, 0, reinterpret_cast<uint32_t*>(pos_w), off);
  env_tables<D, OutT>(cx, v, pl, 0);
  stream_tiles<D, OutT>(cx, P, off, slice, split, ring, out, one16);
}

template <int D, typename OutT>
__global__ void __launch_bounds__(PerspCfg<D>::kWarps * 32)
persp_write_kernel(const uint32_t* __restrict__ syn, const int32_t* __restrict__ local_off,
                   const int64_t* __restrict__ block_off, int epb, OutT* __restrict__ out,
                   uint8_t* __restrict__ pos, int64_t* __restrict__ offsets, int32_t* __restrict__ overflow,
                   int64_t cap, int64_t n, uint32_t one16, int split) {
  using Cfg = PerspCfg<D>;
  constexpr uint32_t kFull = 0xffffffffu;
  extern __shared__ __align__(128) unsigned char smem[];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const WriterCtx<D, OutT> cx(smem, warp, lane);
  build_cand_table<D>(smem, Cfg::kWarps * 32);
  pdl_wait();                 // offsets / syndromes come from the kernels in front; the table above does not
  pdl_launch_dependents();

  uint32_t ring = 0;                             // staging tiles pushed so far by this warp
  const int64_t stride = (int64_t)gridDim.x * Cfg::kWarps;
  // work item = (env, slice): with split > 1 the tiles of one env are divided among `split` warps (large
  // lattices: an env is hundreds of tiles, and few long CTAs would leave the last wave half empty); slice 0
  // also writes the env's positions and offset
  const int64_t n_items = n * split;
  for (int64_t item = (int64_t)blockIdx.x * Cfg::kWarps + warp; item < n_items; item += stride) {
    const int64_t e = split == 1 ? item : item / split;
    const int slice = split == 1 ? 0 : (int)(item - e * split);
    uint32_t v = 0, pl = 0;
    if (lane < D) {
      v = __ldg(syn + e * 2 * D + lane);
      pl = __ldg(syn + e * 2 * D + D + lane);
    }
    int64_t off = 0;
    if (lane == 0) off = __ldg(block_off + e / Cfg::kEpb) + (int64_t)__ldg(local_off + e);   // a shift
    uint32_t m0, m1;
    persp_masks<32>(kFull, D, lane, v, pl, m0, m1);
    const int P = gsum<32>(kFull, __popc(m0) + __popc(m1));
    off = __shfl_sync(kFull, off, 0);
    if (lane == 0 && slice == 0) offsets[e] = off;
    if (P == 0) continue;
    if (off + P > cap) {
      if (lane == 0 && overflow) *overflow = 1;
      continue;
    }
    write_env<D, OutT>(cx, v, pl, m0, m1, P, off, slice, split, ring, out, slice == 0 ? pos : nullptr, one16);
  }
  __syncwarp();
  // the staging tiles must stay valid until the TMA engine has READ them; the global writes themselves complete
  // by grid end like any other store (same as a CUTLASS TMA-store epilogue's tma_store_wait<0>)
  if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// Small lattices (PerspCfg<D>::kGroup > 1): one warp writes kGroup consecutive envs as one stream (see PerspCfg).
template <int D, typename OutT>
__global__ void __launch_bounds__(PerspCfg<D>::kWarps * 32)
persp_write_group_kernel(const uint32_t* __restrict__ syn, const int32_t* __restrict__ local_off,
                         const int64_t* __restrict__ block_off, OutT* __restrict__ out, uint8_t* __restrict__ pos,
                         int64_t* __restrict__ offsets, int32_t* __restrict__ overflow, int64_t cap, int64_t n,
                         uint32_t one16) {
  using Cfg = PerspCfg<D>;
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr int K = Cfg::kGroup, LPE = Cfg::kLanesPerEnv, R = Cfg::kLoadRounds, SPR = 32 / LPE;   // slots per round
  static_assert(K == R * SPR, "group does not tile the load rounds");
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const WriterCtx<D, OutT> cx(smem, warp, lane);
  build_cand_table<D>(smem, Cfg::kWarps * 32);
  pdl_wait();
  pdl_launch_dependents();
  uint32_t ring = 0;
  const int64_t n_items = (n + K - 1) / K;
  const int64_t stride = (int64_t)gridDim.x * Cfg::kWarps;
  for (int64_t item = (int64_t)blockIdx.x * Cfg::kWarps + warp; item < n_items; item += stride) {
    const int64_t e0 = item * K;
    // all loads first: lane = slot * LPE + row holds rows of env e0 + round * SPR + slot
    uint32_t v_all[R], pl_all[R];
#pragma unroll
    for (int rd = 0; rd < R; ++rd) {
      const int64_t e = e0 + rd * SPR + lane / LPE;
      const int row = lane % LPE;
      v_all[rd] = pl_all[rd] = 0;
      if (e < n && row < D) {
        v_all[rd] = __ldg(syn + e * 2 * D + row);
        pl_all[rd] = __ldg(syn + e * 2 * D + D + row);
      }
    }
    int64_t off = 0;
    if (lane == 0) off = __ldg(block_off + e0 / Cfg::kEpb) + (int64_t)__ldg(local_off + e0);
    off = __shfl_sync(kFull, off, 0);
    int kbase = 0, ktrue = 0;                    // perspectives written / counted so far in this group
    bool full = false;                           // capacity reached: this env and the ones behind it are skipped
#pragma unroll
    for (int s = 0; s < K; ++s) {
      const int src = (s % SPR) * LPE + (lane < LPE ? lane : 0);
      uint32_t v = __shfl_sync(kFull, v_all[s / SPR], src);
      uint32_t pl = __shfl_sync(kFull, pl_all[s / SPR], src);
      if (lane >= D) v = pl = 0;
      if (e0 + s >= n) break;                    // warp-uniform
      uint32_t m0, m1;
      persp_masks<32>(kFull, D, lane, v, pl, m0, m1);
      const int P = __reduce_add_sync(kFull, __popc(m0) + __popc(m1));      // redux.sync: one instruction
      if (lane == 0) offsets[e0 + s] = off + ktrue;
      ktrue += P;
      if (P == 0) continue;
      if (full || off + ktrue > cap) {           // offsets grow with the env index: once full, always full
        if (lane == 0 && overflow) *overflow = 1;
        full = true;
        continue;
      }
      env_positions<D, OutT>(cx, m0, m1, kbase, s, reinterpret_cast<uint32_t*>(pos), off);
      env_tables<D, OutT>(cx, v, pl, s);
      kbase += P;
    }
    if (kbase > 0) stream_tiles<D, OutT>(cx, kbase, off, 0, 1, ring, out, one16);
  }
  __syncwarp();
  if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

}  // namespace tc
