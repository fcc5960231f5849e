// Per-environment device routines shared by the standalone kernels and the fused actor step.
// One group of G lanes per environment; lane r (< d) owns row r of every plane.
#pragma once
#include "common.cuh"

namespace tc {

constexpr int kThreads = 256;   // CTA size of every state kernel

template <int G>
__device__ __forceinline__ void load_qubits(const uint32_t* __restrict__ qubits, int64_t e, int d, int r, Rows& s) {
  if (r < d) {
    const uint32_t* q = qubits + e * 4 * d + r;
    s.x0 = q[0];
    s.x1 = q[d];
    s.z0 = q[2 * d];
    s.z1 = q[3 * d];
  } else {
    s.x0 = s.x1 = s.z0 = s.z1 = 0;
  }
}

template <int G>
__device__ __forceinline__ void store_qubits(uint32_t* __restrict__ qubits, int64_t e, int d, int r, const Rows& s) {
  if (r < d) {
    uint32_t* q = qubits + e * 4 * d + r;
    q[0] = s.x0;
    q[d] = s.x1;
    q[2 * d] = s.z0;
    q[3 * d] = s.z1;
  }
}

template <int G>
__device__ __forceinline__ void load_syn(const uint32_t* __restrict__ syn, int64_t e, int d, int r, Rows& s) {
  if (r < d) {
    s.v = syn[e * 2 * d + r];
    s.pl = syn[e * 2 * d + d + r];
  } else {
    s.v = s.pl = 0;
  }
}

template <int G>
__device__ __forceinline__ void store_syn(uint32_t* __restrict__ syn, int64_t e, int d, int r, const Rows& s) {
  if (r < d) {
    syn[e * 2 * d + r] = s.v;
    syn[e * 2 * d + d + r] = s.pl;
  }
}

// ToricCode.step (ToricCode.py:62-88): Pauli product on the acted qubit, then the syndrome of the new
// error configuration.  The reference recomputes the whole syndrome (:83); flipping one qubit only
// touches the two stabilisers it borders, so the update here is incremental:
//   matrix 0: z-part flips V[i,j], V[i+1,j];  x-part flips Pl[i,j], Pl[i,j-1]
//   matrix 1: z-part flips V[i,j], V[i,j+1];  x-part flips Pl[i,j], Pl[i-1,j]
// Returns false (and changes nothing) for an action outside the lattice.
template <int G>
__device__ __forceinline__ bool apply_action(int d, int r, int m, int i, int j, int op, Rows& s) {
  const bool valid = ((unsigned)m < 2u) && ((unsigned)i < (unsigned)d) && ((unsigned)j < (unsigned)d) &&
                     ((unsigned)op < 4u);
  if (!valid) return false;
  const bool xp = (op == 1) || (op == 2);
  const bool zp = (op == 2) || (op == 3);
  const uint32_t bit = 1u << j;
  if (m == 0) {
    if (r == i) {
      if (xp) { s.x0 ^= bit; s.pl ^= bit | (1u << wrap_dec(j, d)); }
      if (zp) { s.z0 ^= bit; s.v ^= bit; }
    }
    if (zp && r == wrap_inc(i, d)) s.v ^= bit;
  } else {
    if (r == i) {
      if (xp) { s.x1 ^= bit; s.pl ^= bit; }
      if (zp) { s.z1 ^= bit; s.v ^= bit | (1u << wrap_inc(j, d)); }
    }
    if (xp && r == wrap_dec(i, d)) s.pl ^= bit;
  }
  return true;
}

// evalGroundState (ToricCode.py:326-335): False if any whole-plane popcount is odd.
template <int G>
__device__ __forceinline__ bool parity_even(uint32_t gmask, const Rows& s) {
  const uint32_t par = (__popc(s.x0) & 1) | ((__popc(s.x1) & 1) << 1) | ((__popc(s.z0) & 1) << 2) |
                       ((__popc(s.z1) & 1) << 3);
  return gxor<G>(gmask, par) == 0;
}

struct SampleStats {
  int attempts;
  int nx, ny, nz;
};

// ToricCode.reset's sampling loop (ToricCode.py:108-117): draw errors until the syndrome is non-trivial.
// Draw numbering (the oracle restates exactly this; see DESIGN.md "RNG"):
//   depolarizing: attempt a, matrix m, row i, column pair k -> Philox counter word3 = a<<10 | m<<9 | i<<4 | k;
//                 output words (u0, c0, u1, c1) serve columns 2k and 2k+1: error iff u < p_thresh,
//                 Pauli code 1 + mulhi(c, 3)                                  (ToricCode.py:141-149)
//   fixed-weight: draw t of attempt a -> word3 = kDomFixed | (a & 1023)<<20 | (t & 0xfffff);
//                 qubit index = u0 & (pow2 - 1), rejected when >= 2d^2 or already occupied; code as above
//                                                                            (ToricCode.py:159-164)
template <int G>
__device__ __forceinline__ void sample_errors(uint32_t gmask, int d, int r, uint64_t env_id, uint32_t episode,
                                              uint32_t thresh, int mode, int n_fixed, uint2 key, Rows& s,
                                              uint8_t& status, SampleStats& st) {
  const bool one_draw = (mode & kSampleOneDraw) != 0;   // the bare generators: no non-trivial-syndrome rejection
  mode &= 3;
  const int nq = 2 * d * d;
  const int nf = n_fixed < nq ? n_fixed : nq;
  uint32_t pm = 1;
  while ((int)pm < nq) pm <<= 1;
  pm -= 1;
  const uint32_t elo = (uint32_t)env_id, ehi = (uint32_t)(env_id >> 32);
  const uint32_t cap = (mode == 0) ? kMaxAttempts : kMaxFixedAttempts;
  uint32_t attempt = 0;
  while (true) {
    s.x0 = s.x1 = s.z0 = s.z1 = 0;
    if (mode != 0) {
      int placed = 0;
      uint32_t t = 0;
      while (placed < nf) {
        const uint4 rnd = philox4x32_10(
            make_uint4(elo, ehi, episode, kDomFixed | ((attempt & 1023u) << 20) | (t & 0xfffffu)), key);
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
    if (mode != 1 && r < d) {
      const uint32_t occ0 = s.x0 | s.z0, occ1 = s.x1 | s.z1;
      const int pairs = (d + 1) >> 1;
#pragma unroll 1
      for (int m = 0; m < 2; ++m) {
        const uint32_t occ = m ? occ1 : occ0;
        uint32_t xr = 0, zr = 0;
        for (int k = 0; k < pairs; ++k) {
          const uint4 rnd = philox4x32_10(
              make_uint4(elo, ehi, episode, (attempt << 10) | ((uint32_t)m << 9) | ((uint32_t)r << 4) | (uint32_t)k),
              key);
          const int j0 = 2 * k, j1 = 2 * k + 1;
          if (rnd.x < thresh && !((occ >> j0) & 1u)) {
            const uint32_t code = 1u + mulhi32(rnd.y, 3u);
            if (code <= 2u) xr |= 1u << j0;
            if (code >= 2u) zr |= 1u << j0;
          }
          if (j1 < d && rnd.z < thresh && !((occ >> j1) & 1u)) {
            const uint32_t code = 1u + mulhi32(rnd.w, 3u);
            if (code <= 2u) xr |= 1u << j1;
            if (code >= 2u) zr |= 1u << j1;
          }
        }
        if (m == 0) { s.x0 |= xr; s.z0 |= zr; } else { s.x1 |= xr; s.z1 |= zr; }
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

// position of the n-th (0-based) set bit of m; m must have more than n bits set
__device__ __forceinline__ int nth_set_bit(uint32_t m, int n) {
  int pos = 0;
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const int c = __popc(m & ((1u << w) - 1u));
    if (n >= c) {
      n -= c;
      pos += w;
      m >>= w;
    }
  }
  return pos;
}

// Uniform pick from the env's perspective list x {X,Y,Z} (the random branch of
// _selectActionBatch_prime, src/numba/util_actor.py:96-98): k = mulhi(u0, P) indexes the list in
// reference order (matrix 0 row-major, then matrix 1), op = 1 + mulhi(u1, 3).
template <int G>
__device__ __forceinline__ void random_action(uint32_t gmask, int d, int r, uint32_t m0, uint32_t m1,
                                              uint64_t env_id, uint32_t episode, uint32_t step, uint2 key,
                                              int& am, int& ai, int& aj, int& aop) {
  const int c0 = __popc(m0), c1 = __popc(m1);
  int t0, t1;
  const int b0 = gscan_excl<G>(gmask, c0, r, t0);
  const int b1 = gscan_excl<G>(gmask, c1, r, t1) + t0;
  const int P = t0 + t1;
  const uint4 rnd = philox4x32_10(make_uint4((uint32_t)env_id, (uint32_t)(env_id >> 32), episode, kDomAction | step), key);
  aop = 1 + (int)mulhi32(rnd.y, 3u);
  int packed = 0;   // m<<16 | i<<8 | j ; stays 0 -> (0,0,0) when the list is empty
  if (P > 0) {
    const int k = (int)mulhi32(rnd.x, (uint32_t)P);
    const bool in0 = k >= b0 && k < b0 + c0, in1 = k >= b1 && k < b1 + c1;
    if (in0 || in1) packed = ((in1 ? 1 : 0) << 16) | (r << 8) | nth_set_bit(in1 ? m1 : m0, in1 ? k - b1 : k - b0);
  }
  // exactly one lane holds the pick: OR-reduce
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) packed |= __shfl_xor_sync(gmask, packed, o, G);
  am = packed >> 16;
  ai = (packed >> 8) & 0xff;
  aj = packed & 0xff;
}

}  // namespace tc
