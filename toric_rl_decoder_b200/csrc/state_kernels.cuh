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

// K1 sampler (sample_errors, SampleStats): sampler.cuh, included at the end of this file

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

#include "sampler.cuh"
