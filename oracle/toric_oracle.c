/* CPU oracle for the toric-code environment hot path -- plain C restatement.
 *
 * TEST INFRASTRUCTURE ONLY.  Built into oracle/libtoric_oracle.so by oracle/Makefile; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and only
 * as the checker or the timed CPU arm.  Nothing in toric_rl_decoder_b200/ links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function here against the golden
 * vectors generated from the unmodified reference (oracle/gen_golden.py) -- the reference is pure
 * Python/numba, so there is nothing to compile into oracle/_ref.
 *
 * The algorithms follow the reference's byte-per-cell formulation (one uint8 per lattice site, like its
 * int arrays), not the bit-packed device layout.  file:line citations are relative to the reference
 * checkout.  The sampler (to_reset / to_random_action) restates the DEVICE's Philox draw numbering
 * (DESIGN.md "RNG"): the reference uses numpy's MT19937, which a counter-based device generator cannot
 * and need not reproduce; its distribution is validated statistically instead.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define IDX(m, i, j) (((m) * d + (i)) * d + (j))
static inline int modd(int a, int d) { a %= d; return a < 0 ? a + d : a; }

/* createSyndromOpt (ToricCode.py:168-199)
 *   V[i,j]  = z0[i,j] ^ z0[i-1,j] ^ z1[i,j] ^ z1[i,j-1]     (:180-186)
 *   Pl[i,j] = x0[i,j] ^ x0[i,j+1] ^ x1[i,j] ^ x1[i+1,j]     (:191-197) */
void to_syndrome(const uint8_t* q, uint8_t* s, int d) {
  for (int i = 0; i < d; ++i)
    for (int j = 0; j < d; ++j) {
      const int im = modd(i - 1, d), ip = modd(i + 1, d), jm = modd(j - 1, d), jp = modd(j + 1, d);
#define ZP(c) ((c) >= 2)
#define XP(c) ((c) == 1 || (c) == 2)
      s[IDX(0, i, j)] = (uint8_t)(ZP(q[IDX(0, i, j)]) ^ ZP(q[IDX(0, im, j)]) ^ ZP(q[IDX(1, i, j)]) ^ ZP(q[IDX(1, i, jm)]));
      s[IDX(1, i, j)] = (uint8_t)(XP(q[IDX(0, i, j)]) ^ XP(q[IDX(0, i, jp)]) ^ XP(q[IDX(1, i, j)]) ^ XP(q[IDX(1, ip, j)]));
    }
}

static int defects(const uint8_t* s, int d) {
  int c = 0;
  for (int k = 0; k < 2 * d * d; ++k) c += s[k];
  return c;
}

/* ToricCode.step + getReward + isTerminalState (ToricCode.py:62-88, :202-218).  Like the reference the
 * whole syndrome is recomputed.  Returns terminal. */
int to_step(uint8_t* q, uint8_t* s, const int32_t* action, int d, double* reward) {
  const int m = action[0], i = action[1], j = action[2], op = action[3];
  const int before = defects(s, d);
  q[IDX(m, i, j)] ^= (uint8_t)op;     /* rule_table[c][op] == c ^ op (ToricCode.py:57) */
  to_syndrome(q, s, d);
  const int after = defects(s, d);
  *reward = after == 0 ? 100.0 : (double)(before - after);
  return after == 0;
}

/* evalGroundState (ToricCode.py:289-337): sticky False when any whole-plane count is odd. */
int to_ground_state(const uint8_t* q, int d, int prev) {
  int px[2] = {0, 0}, pz[2] = {0, 0};
  for (int m = 0; m < 2; ++m)
    for (int k = 0; k < d * d; ++k) {
      const uint8_t c = q[m * d * d + k];
      px[m] += XP(c);
      pz[m] += ZP(c);
    }
  if ((px[0] | px[1] | pz[0] | pz[1]) & 1) return 0;
  return prev;
}

/* generatePerspective (src/util.py:46-85) + rotate_state (:87-94), written straight to float32 like the
 * concatenate/cast of selectActionBatch (src/numba/util_actor.py:37-39).
 * out: P*2*d*d floats; pos: P*3 ints; returns P.  Either output may be NULL (count only). */
int64_t to_perspectives(const uint8_t* s, int d, float* out, int32_t* pos) {
  const int g = d / 2;
  int64_t P = 0;
  for (int m = 0; m < 2; ++m)
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) {
        int sel;
        if (m == 0)   /* src/util.py:68-69 */
          sel = s[IDX(0, i, j)] | s[IDX(0, modd(i + 1, d), j)] | s[IDX(1, i, j)] | s[IDX(1, i, modd(j - 1, d))];
        else          /* src/util.py:77-78 */
          sel = s[IDX(0, i, j)] | s[IDX(0, i, modd(j + 1, d))] | s[IDX(1, i, j)] | s[IDX(1, modd(i - 1, d), j)];
        if (!sel) continue;
        if (pos) { pos[P * 3] = m; pos[P * 3 + 1] = i; pos[P * 3 + 2] = j; }
        if (out) {
          float* o = out + P * 2 * d * d;
          for (int a = 0; a < d; ++a)
            for (int b = 0; b < d; ++b) {
              if (m == 0) {   /* np.roll by (g-i, g-j), src/util.py:70-71 */
                const int r = modd(a + i - g, d), c = modd(b + j - g, d);
                o[IDX(0, a, b)] = (float)s[IDX(0, r, c)];
                o[IDX(1, a, b)] = (float)s[IDX(1, r, c)];
              } else {        /* roll, rot90 CCW, vertex plane rolled +1 on axis 0 (src/util.py:79-81, :87-94) */
                const int r = modd(b + i - g, d);
                o[IDX(0, a, b)] = (float)s[IDX(0, r, modd(modd(d - a, d) + j - g, d))];
                o[IDX(1, a, b)] = (float)s[IDX(1, r, modd(d - 1 - a + j - g, d))];
              }
            }
        }
        ++P;
      }
  return P;
}

/* ------------------------------------------------------------------------------------------------
 * Device sampler spec (Philox4x32-10, Salmon et al. SC'11)
 * ------------------------------------------------------------------------------------------------ */
void to_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline uint32_t mulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

#define DOM_FIXED 0x80000000u
#define DOM_ACTION 0xC0000000u

#define DOM_FIRST 0x40000000u

/* Depolarizing field of one lattice row, both matrices (csrc/sampler.cuh:depolarize_row).  Qubit s = m*d + j of
 * the row owns the 16-bit slot s (bits [16*(s%8), +16) of call s/8); error iff the lazily extended 32-bit uniform
 * is below thresh; the codes of the row's errors are 2-bit chunks (0 rejected) from the stream behind the last
 * slot.  first_q >= 0: linear index q = m*d*d + i*d + j of the forced first error (qubits before it stay clean). */
static void depolarize_row(uint8_t* field /* [2][d][d], written for row `row` only */, int d, int row,
                           uint64_t env_id, uint32_t episode, uint32_t attempt, uint32_t thresh,
                           const uint32_t key[2], int first_q) {
  const uint32_t t_hi = thresh >> 16, t_lo = thresh & 0xffffu;
  const uint32_t base = (attempt << 10) | ((uint32_t)row << 5);
  const int nslots = 2 * d, ncalls = (nslots + 7) / 8;
  uint8_t err[64];
  uint32_t rnd[4] = {0, 0, 0, 0};
  for (int s = 0; s < nslots; ++s) {
    if ((s & 7) == 0) {
      const uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode, base | (uint32_t)(s >> 3)};
      to_philox4x32_10(ctr, key, rnd);
    }
    const int h = s & 7;
    const uint32_t w = rnd[h >> 1];
    const uint32_t u = (h & 1) ? (w >> 16) : (w & 0xffffu);
    err[s] = (uint8_t)(u < t_hi);
    if (u == t_hi) {                              /* draw the low half of the uniform */
      const uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode, base | (uint32_t)(16 + (s >> 2))};
      uint32_t r2[4];
      to_philox4x32_10(ctr, key, r2);
      err[s] = (uint8_t)((r2[s & 3] & 0xffffu) < t_lo);
    }
  }
  if (first_q >= 0)
    for (int s = 0; s < nslots; ++s) {
      const int m = s / d, j = s - m * d, q = (m * d + row) * d + j;
      if (q < first_q) err[s] = 0;
      if (q == first_q) err[s] = 1;
    }
  /* Pauli codes: 2-bit chunks of the stream behind the last slot (word-aligned), chunk 0 rejected.  Slot s first reads
   * chunk s; the error slots whose chunk was 0 then take the chunks from index nslots on, in increasing s. */
  int c = ncalls - 1, wi = (nslots & 7) ? ((nslots & 7) >> 1) : 4, any = 0;
  uint8_t code[64];
  for (int s = 0; s < nslots; ++s) { code[s] = 0; any |= err[s]; }
  if (any) {
    uint32_t words[4] = {0, 0, 0, 0}, cur = 0;
    const int nwords = (nslots + 15) / 16;
    int left;
#define NEXT_WORD(dst)                                                                                             \
  do {                                                                                                             \
    if (wi == 4) {                                                                                                 \
      ++c;                                                                                                         \
      const uint32_t ctr_[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode, base | (uint32_t)(c & 15)};   \
      to_philox4x32_10(ctr_, key, rnd);                                                                            \
      wi = 0;                                                                                                      \
    }                                                                                                              \
    (dst) = rnd[wi++];                                                                                             \
  } while (0)
    for (int j = 0; j < nwords; ++j) NEXT_WORD(words[j]);
    for (int s = 0; s < nslots; ++s)
      if (err[s]) code[s] = (uint8_t)((words[s >> 4] >> (2 * (s & 15))) & 3u);
    left = (16 - (nslots & 15)) & 15;
    cur = left ? (words[nwords - 1] >> (2 * (16 - left))) : 0u;
    for (int s = 0; s < nslots; ++s) {
      if (!err[s] || code[s]) continue;
      uint32_t v;
      do {
        if (left == 0) { NEXT_WORD(cur); left = 16; }
        v = cur & 3u;
        cur >>= 2;
        --left;
      } while (v == 0);
      code[s] = (uint8_t)v;
    }
#undef NEXT_WORD
  }
  for (int s = 0; s < nslots; ++s) {
    const int m = s / d, j = s - m * d;
    field[IDX(m, row, j)] = code[s];              /* 0 = no error, else the Pauli code */
  }
}

/* ToricCode.reset's sampling loop (ToricCode.py:108-117) with the device's draw numbering (csrc/sampler.cuh).
 * mode 0: generateRandomError (:130-151)   mode 1: generateNRandomErrors (:153-166), n_fixed distinct qubits
 * mode 2: generateNPlusQRandomErrors (results/small_p_error_test.py:43-52);  mode | 16: one draw, no rejection.
 * When thresh * 2d^2 < 2^32 mode 0 draws the field conditioned on >= 1 error (first-error index from a
 * truncated geometric law), which leaves the distribution after the rejection unchanged.
 * Returns the number of attempts (0x80000000 | attempts if the sampler gave up). */
uint32_t to_reset(uint8_t* q, uint8_t* s, int d, uint64_t env_id, uint32_t episode, uint32_t thresh, int mode,
                  int n_fixed, uint64_t seed) {
  const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  const int one_draw = mode & 16;     /* bare generator call: no rejection of trivial syndromes */
  mode &= 3;
  const int nq = 2 * d * d;
  const int nf = n_fixed < nq ? n_fixed : nq;
  uint32_t pm = 1;
  while ((int)pm < nq) pm <<= 1;
  pm -= 1;
  const uint32_t cap = mode == 0 ? (1u << 20) : (1u << 9);   /* 9 attempt bits in the fixed-weight counter (common.cuh) */
  const int conditioned = mode == 0 && !one_draw && (uint64_t)thresh * (uint64_t)nq < (8ull << 32);
  uint32_t attempt = 0, rnd[4];
  uint8_t field[2 * 32 * 32];
  if (conditioned && thresh == 0) {
    memset(q, 0, (size_t)nq);
    memset(s, 0, (size_t)nq);
    return 0x80000000u | 1u;
  }
  double log_q = 0.0, mass = 0.0;
  if (conditioned) {
    log_q = log1p(-(double)thresh * (1.0 / 4294967296.0));
    mass = -expm1((double)nq * log_q);
  }
  for (;;) {
    memset(q, 0, (size_t)nq);
    if (mode != 0) {
      int placed = 0;
      uint32_t t = 0;
      while (placed < nf) {
        const uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode,
                                 DOM_FIXED | ((attempt & 511u) << 20) | (t & 0xfffffu)};
        to_philox4x32_10(ctr, key, rnd);
        ++t;
        const int qi = (int)(rnd[0] & pm);
        if (qi >= nq || q[qi]) continue;
        q[qi] = (uint8_t)(1u + mulhi(rnd[1], 3u));
        ++placed;
      }
    }
    if (mode != 1) {
      int first_q = -1;
      if (conditioned) {
        const uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode, DOM_FIRST | attempt};
        to_philox4x32_10(ctr, key, rnd);
        const uint64_t u53 = (((uint64_t)rnd[0] << 32) | (uint64_t)rnd[1]) >> 11;
        const double u = (double)u53 * (1.0 / 9007199254740992.0);
        const double k = floor(log1p(-u * mass) / log_q);
        first_q = k >= (double)(nq - 1) ? nq - 1 : (k > 0.0 ? (int)k : 0);
      }
      for (int i = 0; i < d; ++i) depolarize_row(field, d, i, env_id, episode, attempt, thresh, key, first_q);
      for (int k = 0; k < nq; ++k)
        if (!q[k]) q[k] = field[k];               /* occupied qubits keep the fixed part's code */
    }
    to_syndrome(q, s, d);
    ++attempt;
    if (defects(s, d) != 0 || one_draw) return attempt;
    if (attempt >= cap) return 0x80000000u | attempt;
  }
}

/* to_reset over a batch of envs (ids env_base .. env_base+n-1, same episode / thresh); q, s: uint8 [n][2][d][d];
 * attempts (nullable): uint32 [n] return values of to_reset. */
void to_reset_batch(uint8_t* q, uint8_t* s, int d, int64_t n, uint64_t env_base, uint32_t episode, uint32_t thresh,
                    int mode, int n_fixed, uint64_t seed, uint32_t* attempts) {
  const int nq = 2 * d * d;
  for (int64_t e = 0; e < n; ++e) {
    const uint32_t a = to_reset(q + e * nq, s + e * nq, d, env_base + (uint64_t)e, episode, thresh, mode, n_fixed, seed);
    if (attempts) attempts[e] = a;
  }
}

/* Random branch of _selectActionBatch_prime (src/numba/util_actor.py:96-98) with the device's draws:
 * k = mulhi(u0, P) into the perspective list (reference order), op = 1 + mulhi(u1, 3). */
void to_random_action(const uint8_t* s, int d, uint64_t env_id, uint32_t episode, uint32_t step, uint64_t seed,
                      int32_t* action, int32_t* pos_scratch) {
  const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  const uint32_t ctr[4] = {(uint32_t)env_id, (uint32_t)(env_id >> 32), episode, DOM_ACTION | step};
  uint32_t rnd[4];
  to_philox4x32_10(ctr, key, rnd);
  const int64_t P = to_perspectives(s, d, NULL, pos_scratch);
  action[0] = action[1] = action[2] = 0;
  action[3] = 1 + (int)mulhi(rnd[1], 3u);
  if (P > 0) {
    const int k = (int)mulhi(rnd[0], (uint32_t)P);
    action[0] = pos_scratch[k * 3];
    action[1] = pos_scratch[k * 3 + 1];
    action[2] = pos_scratch[k * 3 + 2];
  }
}

/* ------------------------------------------------------------------------------------------------
 * Batched actor loop = what bench.py times as the CPU arm: the loop body of src/Actor_mp.py:104-185
 * without the network (random policy), over n envs, `steps` times.  Single-threaded like one reference
 * actor process; the caller runs one instance per host core on its own slice of envs (the reference's
 * one-process-per-actor design, Distributed_mp.py:199-211) -- ctypes releases the GIL during the call.
 *   q, s        uint8 [n][2][d][d] qubits / syndromes (in/out)
 *   step_ctr    int32 [n]   steps in the current episode (in/out);  episode uint32 [n] (in/out)
 *   out         float32, capacity cap perspectives; rewritten every step like the device buffer
 *   counters    int64 [4]: env steps, episodes, solved, perspectives
 * Returns 0, or -1 when cap is too small.
 * ------------------------------------------------------------------------------------------------ */
int to_actor_steps(uint8_t* q, uint8_t* s, int32_t* step_ctr, uint32_t* episode, const uint32_t* thresh,
                   int64_t n, int d, int steps, int max_steps, uint64_t seed, int64_t env_base, float* out,
                   int32_t* pos, int64_t cap, int64_t* offsets, int64_t* counters) {
  const int nq = 2 * d * d;
  int32_t* scratch = (int32_t*)malloc(sizeof(int32_t) * 3 * (size_t)nq);
  int rc = 0;
  for (int t = 0; t < steps && rc == 0; ++t) {
    int64_t episodes = 0, solved = 0;
    for (int64_t e = 0; e < n; ++e) {            /* EnvSet.step's loop (src/EnvSet.py:42-46) */
      uint8_t* qe = q + e * nq;
      uint8_t* se = s + e * nq;
      int32_t act[4];
      double rew;
      to_random_action(se, d, (uint64_t)(env_base + e), episode[e], (uint32_t)step_ctr[e], seed, act, scratch);
      const int term = to_step(qe, se, act, d, &rew);
      step_ctr[e] += 1;
      if (term || step_ctr[e] > max_steps) {     /* Actor_mp.py:171-183 */
        episodes += 1;
        solved += term;
        to_reset(qe, se, d, (uint64_t)(env_base + e), episode[e], thresh[e], 0, 0, seed);
        episode[e] += 1;
        step_ctr[e] = 0;
      }
    }
    /* generatePerspectiveBatch + concatenate + float32 (src/numba/util_actor.py:33-39, :56-67) */
    offsets[0] = 0;
    for (int64_t e = 0; e < n; ++e) {
      const int64_t room = cap - offsets[e];
      if (room < nq) { rc = -1; break; }
      offsets[e + 1] = offsets[e] + to_perspectives(s + e * nq, d, out + offsets[e] * nq,
                                                    pos ? pos + offsets[e] * 3 : NULL);
    }
    counters[0] += n;
    counters[1] += episodes;
    counters[2] += solved;
    counters[3] += offsets[n];
  }
  free(scratch);
  return rc;
}
