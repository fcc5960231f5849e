// K1: the error sampler of ToricCode.reset (gym_ToricCode/.../ToricCode.py:90-166) and of
// generateNPlusQRandomErrors (results/small_p_error_test.py:43-52) on Philox4x32-10.
//
// One group of G lanes per environment; lane i (< d) owns lattice row i of both qubit matrices, i.e. the 2d
// qubits s = m*d + j (matrix 0 columns 0..d-1, then matrix 1).  The whole sampler stays in registers.
//
// DRAW NUMBERING (restated bit for bit by oracle/toric_oracle.c:to_reset -- keep the two in step; the Philox
// counter is (env id lo, env id hi, episode, word3), the key is the seed):
//
//  depolarizing field (generateRandomError, ToricCode.py:130-151: error iff U(0,1) < p, Pauli ~ U{1,2,3}):
//    word3 = attempt<<10 | row<<5 | c.  Calls c = 0,1,.. are 128-bit blocks (w0 | w1<<32 | w2<<64 | w3<<96).
//    * qubit s of the row owns the 16-bit slot s: bits [16*(s%8), +16) of call s/8 = u16.  With
//      T = floor(p * 2^32):  u16 < T>>16 -> error;  u16 > T>>16 -> no error;  u16 == T>>16 (probability
//      2^-16) -> refined by word s%4 of call 16 + s/4:  error iff (word & 0xffff) < (T & 0xffff).
//      Together that is exactly "a 32-bit uniform < T", the uniform's low half being drawn only when needed --
//      8 qubits per Philox call instead of 2.
//    * Pauli codes: 2-bit chunks of the bit stream that starts right behind the row's last slot (bit 16*(2d%8) of call
//      (2d-1)/8; call 2d/8 when 2d%8 == 0) and runs on through the following calls (c & 15).  Chunk 0 is rejected, chunk
//      v in {1,2,3} is the Pauli code -- exactly uniform.  Slot s first reads chunk s of the stream (positional, so the
//      whole row is assigned bit-parallel); the error slots whose chunk was 0, in increasing s, then take the chunks from
//      index 2d on, one after the other, until theirs is non-zero.
//  rejection (ToricCode.py:108-117): attempts are repeated until the syndrome is non-trivial.  When
//    T * 2d^2 < 8 * 2^32 (p * #qubits < 8: an empty draw is not yet negligible, e^-8) the field is drawn CONDITIONED
//    on holding at least one error -- same distribution after the rejection, O(1) attempts at any p:
//    word3 = 0x40000000 | attempt gives U = ((w0<<32 | w1) >> 11) * 2^-53; the first error sits at linear index
//    K = min(2d^2-1, floor(log1p(-U*A) / log1p(-p))), A = 1 - (1-p)^(2d^2), in the order q = m*d*d + i*d + j
//    (truncated geometric law); qubits q < K are error-free, q == K errs, q > K follow the field above
//    (slots are positional: skipped qubits simply leave theirs unused).  p == 0 has no such draw: the env is left
//    error-free and flagged kStatusResetGaveUp.
//  fixed-weight part (generateNRandomErrors, ToricCode.py:153-166): draw t of attempt a ->
//    word3 = 0x80000000 | (a & 511)<<20 | (t & 0xfffff); qubit index = w0 & (pow2 - 1), rejected when >= 2d^2 or
//    already occupied; code = 1 + mulhi(w1, 3).  Mode 2 then overlays the depolarizing field on the qubits
//    that are still free (occupied qubits keep their slot and, if it says "error", their code chunks).
#pragma once
#include <type_traits>

#include "common.cuh"

namespace tc {

struct SampleStats {
  int attempts;
  int nx, ny, nz;
};

__device__ __forceinline__ uint32_t word_of(const uint4& v, int k) {
  return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
}

// Index of the first error of a depolarizing field conditioned on holding at least one: truncated geometric law
// P(K = k) = (1-p)^k p / (1 - (1-p)^nq), drawn by inversion from a 53-bit uniform (w0, w1 of the kDomFirst call).
// Not inlined: the double-precision logarithms would otherwise cost the whole sampler ~25 registers of occupancy for a
// path only low-p resets take.
__device__ __noinline__ int first_error_index(uint32_t thresh, int nq, uint32_t w0, uint32_t w1) {
  const double log_q = log1p(-(double)thresh * (1.0 / 4294967296.0));        // log(1 - p)
  const double mass = -expm1((double)nq * log_q);                            // 1 - (1-p)^nq
  const unsigned long long u53 = (((unsigned long long)w0 << 32) | (unsigned long long)w1) >> 11;
  const double u = (double)u53 * (1.0 / 9007199254740992.0);
  const double k = floor(log1p(-u * mass) / log_q);
  return k >= (double)(nq - 1) ? nq - 1 : (k > 0.0 ? (int)k : 0);
}

// bits 0, 2, 4, .. 30 of x packed into bits 0..15
__device__ __forceinline__ uint32_t even_bits(uint32_t x) {
  x &= 0x55555555u;
  x = (x | (x >> 1)) & 0x33333333u;
  x = (x | (x >> 2)) & 0x0f0f0f0fu;
  x = (x | (x >> 4)) & 0x00ff00ffu;
  x = (x | (x >> 8)) & 0x0000ffffu;
  return x;
}

__device__ __forceinline__ int first_set(uint32_t m) { return __ffs((int)m) - 1; }
__device__ __forceinline__ int first_set(unsigned long long m) { return __ffsll((long long)m) - 1; }

// Depolarizing field of one lattice row (both matrices).  first_q: linear index of the forced first error, or
// -1 (unconditioned).  Returns the row's x / z planes of matrix 0 and 1.  M: the row's slot mask type -- uint32_t while
// 2d <= 32 (group sizes 8 and 16), unsigned long long above (half the instructions of the 64-bit form where it fits).
template <typename M>
__device__ __forceinline__ void depolarize_row(int d, int row, uint32_t elo, uint32_t ehi, uint32_t episode,
                                               uint32_t attempt, uint32_t thresh, uint2 key, int first_q,
                                               uint32_t& x0, uint32_t& z0, uint32_t& x1, uint32_t& z1) {
  constexpr int kBits = (int)sizeof(M) * 8;
  const uint32_t t_hi = thresh >> 16, t_lo = thresh & 0xffffu;
  const uint32_t base = (attempt << 10) | ((uint32_t)row << 5);
  const int nslots = 2 * d;
  const int ncalls = (nslots + 7) >> 3;
  M err = 0, bnd = 0;                            // bit s = qubit s of the row
  uint4 rnd = make_uint4(0, 0, 0, 0);
#pragma unroll 1
  for (int c = 0; c < ncalls; ++c) {
    rnd = philox4x32_10(make_uint4(elo, ehi, episode, base | (uint32_t)c), key);
    uint32_t lt = 0, eq = 0;
#pragma unroll
    for (int h = 0; h < 8; ++h) {
      const uint32_t w = word_of(rnd, h >> 1);
      const uint32_t u = (h & 1) ? (w >> 16) : (w & 0xffffu);
      lt |= (uint32_t)(u < t_hi) << h;
      eq |= (uint32_t)(u == t_hi) << h;
    }
    err |= (M)lt << (8 * c);
    bnd |= (M)eq << (8 * c);
  }
  const M live = nslots >= kBits ? ~(M)0 : (((M)1 << nslots) - (M)1);
  err &= live;
  bnd &= live;
  while (bnd) {                                  // probability 2^-16 per qubit: draw the uniform's low half
    const int s = first_set(bnd);
    bnd &= bnd - 1;
    const uint4 r = philox4x32_10(make_uint4(elo, ehi, episode, base | (uint32_t)(16 + (s >> 2))), key);
    if ((word_of(r, s & 3) & 0xffffu) < t_lo) err |= (M)1 << s;
  }
  if (first_q >= 0) {                            // qubits before the first error are clean, the first one errs
    const int dd = d * d;
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int rel = first_q - (m * dd + row * d);        // column of the first error inside this row, if any
      const M clean = rel <= 0 ? (M)0 : (rel >= d ? (((M)1 << d) - (M)1) : (((M)1 << rel) - (M)1));
      const M one = (rel >= 0 && rel < d) ? ((M)1 << rel) : (M)0;
      err = (err & ~(clean << (m * d))) | (one << (m * d));
    }
  }
  // Pauli codes: 2-bit chunks (chunk 0 rejected, chunk v in {1,2,3} = the code) from the words behind the last slot --
  // 2d is even, so the stream starts on a word boundary: word (2d % 8) / 2 of the last call, or the next call when that
  // one is full.  First pass, bit-parallel: slot s reads chunk s (low bit -> plane A, high bit -> plane B); second pass:
  // the error slots whose chunk was 0 (a quarter of them), in increasing s, take further chunks from chunk 2d on.
  int c = ncalls - 1;
  int wi = (nslots & 7) ? ((nslots & 7) >> 1) : 4;   // next word of the current call (4 = exhausted)
  auto next_word = [&]() -> uint32_t {
    if (wi == 4) {
      ++c;
      rnd = philox4x32_10(make_uint4(elo, ehi, episode, base | (uint32_t)(c & 15)), key);
      wi = 0;
    }
    return word_of(rnd, wi++);
  };
  M pa = 0, pb = 0;
  uint32_t cur = 0;
  int left = 0;
  if (err) {                                     // a row without errors reads nothing
    const int nwords = (nslots + 15) >> 4;
#pragma unroll 1
    for (int j = 0; j < nwords; ++j) {
      cur = next_word();
      pa |= (M)even_bits(cur) << (16 * j);
      pb |= (M)even_bits(cur >> 1) << (16 * j);
    }
    left = (16 - (nslots & 15)) & 15;            // chunks left in `cur` behind chunk 2d - 1
    cur = left ? (cur >> (2 * (16 - left))) : 0u;
  }
  M xs = err & (pa ^ pb);                        // 1:X 2:Y 3:Z -> x = code in {1,2}, z = code in {2,3}
  M zs = err & pb;
  M todo = err & ~(pa | pb);
  while (todo) {
    const int s = first_set(todo);
    todo &= todo - 1;
    uint32_t v;
    do {
      if (left == 0) {
        cur = next_word();
        left = 16;
      }
      v = cur & 3u;
      cur >>= 2;
      --left;
    } while (v == 0u);
    if (v != 3u) xs |= (M)1 << s;
    if (v != 1u) zs |= (M)1 << s;
  }
  const uint32_t rm = row_mask(d);
  x0 = (uint32_t)xs & rm;
  z0 = (uint32_t)zs & rm;
  x1 = (uint32_t)(xs >> d) & rm;
  z1 = (uint32_t)(zs >> d) & rm;
}

// ToricCode.reset's sampling loop (ToricCode.py:108-117): draw errors until the syndrome is non-trivial.
template <int G>
__device__ __forceinline__ void sample_errors(uint32_t gmask, int d, int r, uint64_t env_id, uint32_t episode,
                                              uint32_t thresh, int mode, int n_fixed, uint2 key, Rows& s,
                                              uint8_t& status, SampleStats& st) {
  const bool one_draw = (mode & kSampleOneDraw) != 0;   // the bare generators: no non-trivial-syndrome rejection
  mode &= 3;
  const int nq = 2 * d * d;
  const int nf = n_fixed < nq ? n_fixed : nq;
  const uint32_t pm = 0xffffffffu >> __clz(nq - 1);      // smallest 2^k - 1 >= nq - 1 (nq >= 18)
  const uint32_t elo = (uint32_t)env_id, ehi = (uint32_t)(env_id >> 32);
  const uint32_t cap = (mode == 0) ? kMaxAttempts : kMaxFixedAttempts;
  const bool conditioned = mode == 0 && !one_draw && (unsigned long long)thresh * (unsigned long long)nq < (kCondMeanErrors << 32);
  uint32_t attempt = 0;
  if (conditioned && thresh == 0u) {             // p == 0: nothing to condition on
    s.x0 = s.x1 = s.z0 = s.z1 = 0;
    s.v = s.pl = 0;
    status |= kStatusResetGaveUp;
    st.attempts = 1;
    st.nx = st.ny = st.nz = 0;
    return;
  }
  while (true) {
    s.x0 = s.x1 = s.z0 = s.z1 = 0;
    if (mode != 0) {
      int placed = 0;
      uint32_t t = 0;
      while (placed < nf) {
        const uint4 rnd = philox4x32_10(
            make_uint4(elo, ehi, episode, kDomFixed | ((attempt & 511u) << 20) | (t & 0xfffffu)), key);
        ++t;
        const int q = (int)(rnd.x & pm);
        if (q >= nq) continue;
        const int m = q / (d * d);
        const int rem = q - m * d * d;
        const int i = rem / d;
        const int j = rem - i * d;
        const uint32_t mine = ((m == 0 ? (s.x0 | s.z0) : (s.x1 | s.z1)) >> j) & 1u;
        if (gshfl<G>(gmask, mine, i)) continue;
        const uint32_t code = 1u + mulhi32(rnd.y, 3u);
        if (r == i) {
          const uint32_t bit = 1u << j;
          const uint32_t xb = (code == 1u || code == 2u) ? bit : 0u;
          const uint32_t zb = (code >= 2u) ? bit : 0u;
          if (m == 0) { s.x0 |= xb; s.z0 |= zb; } else { s.x1 |= xb; s.z1 |= zb; }
        }
        ++placed;
      }
    }
    if (mode != 1) {
      int first_q = -1;
      if (conditioned) {
        const uint4 rf = philox4x32_10(make_uint4(elo, ehi, episode, kDomFirst | attempt), key);
        first_q = first_error_index(thresh, nq, rf.x, rf.y);
      }
      if (r < d) {
        const uint32_t occ0 = s.x0 | s.z0, occ1 = s.x1 | s.z1;
        uint32_t x0, z0, x1, z1;
        using Mask = typename std::conditional<(G <= 16), uint32_t, unsigned long long>::type;
        depolarize_row<Mask>(d, r, elo, ehi, episode, attempt, thresh, key, first_q, x0, z0, x1, z1);
        s.x0 |= x0 & ~occ0;
        s.z0 |= z0 & ~occ0;
        s.x1 |= x1 & ~occ1;
        s.z1 |= z1 & ~occ1;
      }
    }
    syndrome_rows<G>(gmask, d, r, s);
    const uint32_t any = __ballot_sync(gmask, (s.v | s.pl) != 0u) & gmask;
    ++attempt;
    if (any || one_draw) break;
    if (attempt >= cap) { status |= kStatusResetGaveUp; break; }
  }
  st.attempts = (int)attempt;
  st.nx = __popc(s.x0 & ~s.z0) + __popc(s.x1 & ~s.z1);
  st.ny = __popc(s.x0 & s.z0) + __popc(s.x1 & s.z1);
  st.nz = __popc(s.z0 & ~s.x0) + __popc(s.z1 & ~s.x1);
}

}  // namespace tc
