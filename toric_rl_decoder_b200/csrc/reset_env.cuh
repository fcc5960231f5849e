// K1 for small lattices (d <= 8), depolarizing mode: ONE THREAD PER ENVIRONMENT.
//
// reset_kernel gives an env a group of 8 lanes (lane = lattice row).  ncu (profiles/r02_ncu_reset_d7*.txt) shows that
// kernel bound by integer instruction issue with 20 of 32 threads active on average: 7 of 8 lanes hold a row, every
// group-uniform quantity is computed 8 times, the Pauli-code loop runs for the slowest row of the warp.  Here a thread
// draws all 2d rows of its env itself (same draw numbering -- csrc/sampler.cuh:depolarize_row -- so the result is
// bit-identical), keeps the four qubit planes as 64-bit words (row r at bits [r*d, r*d + d), d*d <= 64), computes the
// syndrome with four shifts per plane, and the CTA writes the packed rows out through shared memory so that global
// stores stay coalesced although consecutive threads own consecutive envs.
#pragma once
#include "sampler.cuh"

namespace tc {

constexpr int kResetEnvThreads = 128;

template <int D>
__global__ void __launch_bounds__(kResetEnvThreads)
reset_env_kernel(uint32_t* __restrict__ qubits, uint32_t* __restrict__ syn, uint8_t* __restrict__ gs,
                 uint8_t* __restrict__ status, const uint32_t* __restrict__ p_thresh, uint32_t* __restrict__ episode,
                 const int64_t* __restrict__ idx, int64_t n_idx, int one_draw, uint2 key, int64_t env_base) {
  static_assert(D >= kMinD && D <= 8, "one plane must fit 64 bits");
  constexpr int DD = D * D;
  constexpr int nq = 2 * DD;
  constexpr unsigned long long kAll = DD >= 64 ? ~0ull : ((1ull << (DD & 63)) - 1ull);
  constexpr uint32_t kRow = (1u << D) - 1u;
  // staging in the HBM layout itself (env-major, dense): the copy-out is then a linear, coalesced memcpy without any
  // index arithmetic; the thread-major staging stores take a 4-way bank conflict (stride 4d / 2d words), which is cheap
  __shared__ uint32_t stage_q[kResetEnvThreads * 4 * D];
  __shared__ uint32_t stage_s[kResetEnvThreads * 2 * D];
  unsigned long long col0 = 0;                   // column 0 of every row
#pragma unroll
  for (int i = 0; i < D; ++i) col0 |= 1ull << (i * D);
  const unsigned long long colL = col0 << (D - 1);

  const int64_t k0 = (int64_t)blockIdx.x * kResetEnvThreads;
  const int64_t k = k0 + threadIdx.x;
  const bool valid = k < n_idx;
  const int64_t e = valid ? (idx ? idx[k] : k) : 0;
  unsigned long long x0 = 0, x1 = 0, z0 = 0, z1 = 0, v = 0, pl = 0;
  uint32_t ep = 0;
  uint8_t st = 0;
  if (valid) {
    ep = episode[e];
    const uint32_t thresh = p_thresh[e];
    const uint64_t env_id = (uint64_t)(env_base + e);
    const uint32_t elo = (uint32_t)env_id, ehi = (uint32_t)(env_id >> 32);
    const bool conditioned = !one_draw && (unsigned long long)thresh * (unsigned long long)nq < (kCondMeanErrors << 32);
    if (conditioned && thresh == 0u) {
      st = kStatusResetGaveUp;                   // p == 0: nothing to condition on (sampler.cuh)
    } else {
      uint32_t attempt = 0;
      while (true) {
        int first_q = -1;
        if (conditioned) {
          const uint4 rf = philox4x32_10(make_uint4(elo, ehi, ep, kDomFirst | attempt), key);
          first_q = first_error_index(thresh, nq, rf.x, rf.y);
        }
        x0 = x1 = z0 = z1 = 0;
#pragma unroll 1
        for (int row = 0; row < D; ++row) {
          uint32_t rx0, rz0, rx1, rz1;
          depolarize_row<uint32_t>(D, row, elo, ehi, ep, attempt, thresh, key, first_q, rx0, rz0, rx1, rz1);
          const int sh = row * D;
          x0 |= (unsigned long long)rx0 << sh;
          z0 |= (unsigned long long)rz0 << sh;
          x1 |= (unsigned long long)rx1 << sh;
          z1 |= (unsigned long long)rz1 << sh;
        }
        // createSyndromOpt (ToricCode.py:168-199) on whole planes:
        //   V[i,j]  = z0[i,j] ^ z0[i-1,j] ^ z1[i,j] ^ z1[i,j-1];   Pl[i,j] = x0[i,j] ^ x0[i,j+1] ^ x1[i,j] ^ x1[i+1,j]
        const unsigned long long z0_up = ((z0 << D) | (z0 >> (DD - D))) & kAll;               // row i <- row i-1
        const unsigned long long z1_lf = (((z1 << 1) & ~col0) | ((z1 >> (D - 1)) & col0)) & kAll;   // col j <- col j-1
        const unsigned long long x0_rt = ((x0 >> 1) & ~colL) | ((x0 << (D - 1)) & colL);      // col j <- col j+1
        const unsigned long long x1_dn = ((x1 >> D) | (x1 << (DD - D))) & kAll;               // row i <- row i+1
        v = z0 ^ z0_up ^ z1 ^ z1_lf;
        pl = x0 ^ x0_rt ^ x1 ^ x1_dn;
        ++attempt;
        if ((v | pl) != 0ull || one_draw) break;
        if (attempt >= kMaxAttempts) { st = kStatusResetGaveUp; break; }
      }
    }
  }
  // rows of this env in the HBM layout: qubits [4][d], syn [2][d]
  uint32_t* mq = stage_q + threadIdx.x * 4 * D;
  uint32_t* ms = stage_s + threadIdx.x * 2 * D;
#pragma unroll
  for (int r = 0; r < D; ++r) {
    mq[0 * D + r] = (uint32_t)(x0 >> (r * D)) & kRow;
    mq[1 * D + r] = (uint32_t)(x1 >> (r * D)) & kRow;
    mq[2 * D + r] = (uint32_t)(z0 >> (r * D)) & kRow;
    mq[3 * D + r] = (uint32_t)(z1 >> (r * D)) & kRow;
    ms[0 * D + r] = (uint32_t)(v >> (r * D)) & kRow;
    ms[1 * D + r] = (uint32_t)(pl >> (r * D)) & kRow;
  }
  __syncthreads();
  if (idx == nullptr) {                          // the CTA's envs are consecutive: linear coalesced copies
    const int64_t rest = n_idx - k0;
    const int cnt = rest < kResetEnvThreads ? (int)rest : kResetEnvThreads;
    uint32_t* gq = qubits + k0 * 4 * D;
    for (int w = threadIdx.x; w < cnt * 4 * D; w += kResetEnvThreads) gq[w] = stage_q[w];
    uint32_t* gsyn = syn + k0 * 2 * D;
    for (int w = threadIdx.x; w < cnt * 2 * D; w += kResetEnvThreads) gsyn[w] = stage_s[w];
  } else if (valid) {                            // listed envs (resetTerminalEnvs): scattered, written per thread
    uint32_t* gq = qubits + e * 4 * D;
    uint32_t* gsyn = syn + e * 2 * D;
#pragma unroll
    for (int w = 0; w < 4 * D; ++w) gq[w] = mq[w];
#pragma unroll
    for (int w = 0; w < 2 * D; ++w) gsyn[w] = ms[w];
  }
  if (valid) {
    episode[e] = ep + 1;
    if (gs) gs[e] = 1;
    if (status && st) status[e] |= st;
  }
}

// The same idea for 9 <= d <= 16 (a row pair still fits a 32-bit slot mask, a plane no longer fits a register pair): the
// thread keeps its env's 6d row words in shared memory, word-major with a padded pitch -- S(w, t) = smem[w * (T + 1) + t] --
// so that both the thread-major accesses while sampling (consecutive t) and the env-major copy-out (consecutive w of one
// env: bank (w + env) % 32) are free of bank conflicts.
constexpr int kResetRowsThreads = 128;
constexpr int kResetRowsPitch = kResetRowsThreads + 1;

template <int D>
constexpr int reset_rows_smem_bytes() { return 6 * D * kResetRowsPitch * 4; }

template <int D>
__global__ void __launch_bounds__(kResetRowsThreads)
reset_rows_kernel(uint32_t* __restrict__ qubits, uint32_t* __restrict__ syn, uint8_t* __restrict__ gs,
                  uint8_t* __restrict__ status, const uint32_t* __restrict__ p_thresh, uint32_t* __restrict__ episode,
                  const int64_t* __restrict__ idx, int64_t n_idx, int one_draw, uint2 key, int64_t env_base) {
  static_assert(D > 8 && D <= 16, "row pairs must fit a 32-bit slot mask");
  constexpr int nq = 2 * D * D;
  extern __shared__ uint32_t rows_smem[];
  const int t = threadIdx.x;
  uint32_t* S = rows_smem + t;                   // S[w * pitch]: word w of this thread's env
  constexpr int P = kResetRowsPitch;
  const int64_t k0 = (int64_t)blockIdx.x * kResetRowsThreads;
  const int64_t k = k0 + t;
  const bool valid = k < n_idx;
  const int64_t e = valid ? (idx ? idx[k] : k) : 0;
  uint32_t ep = 0;
  uint8_t st = 0;
  if (valid) {
    ep = episode[e];
    const uint32_t thresh = p_thresh[e];
    const uint64_t env_id = (uint64_t)(env_base + e);
    const uint32_t elo = (uint32_t)env_id, ehi = (uint32_t)(env_id >> 32);
    const bool conditioned = !one_draw && (unsigned long long)thresh * (unsigned long long)nq < (kCondMeanErrors << 32);
    if (conditioned && thresh == 0u) {
      st = kStatusResetGaveUp;
#pragma unroll 1
      for (int w = 0; w < 6 * D; ++w) S[w * P] = 0u;
    } else {
      uint32_t attempt = 0;
      while (true) {
        int first_q = -1;
        if (conditioned) {
          const uint4 rf = philox4x32_10(make_uint4(elo, ehi, ep, kDomFirst | attempt), key);
          first_q = first_error_index(thresh, nq, rf.x, rf.y);
        }
#pragma unroll 1
        for (int row = 0; row < D; ++row) {
          uint32_t rx0, rz0, rx1, rz1;
          depolarize_row<uint32_t>(D, row, elo, ehi, ep, attempt, thresh, key, first_q, rx0, rz0, rx1, rz1);
          S[(0 * D + row) * P] = rx0;
          S[(1 * D + row) * P] = rx1;
          S[(2 * D + row) * P] = rz0;
          S[(3 * D + row) * P] = rz1;
        }
        uint32_t any = 0;
#pragma unroll 1
        for (int r = 0; r < D; ++r) {            // createSyndromOpt (ToricCode.py:168-199), row by row
          const int up = r == 0 ? D - 1 : r - 1, dn = r + 1 == D ? 0 : r + 1;
          const uint32_t z1r = S[(3 * D + r) * P], x0r = S[(0 * D + r) * P];
          const uint32_t v = S[(2 * D + r) * P] ^ S[(2 * D + up) * P] ^ z1r ^ rotl_d(z1r, D);
          const uint32_t pl = x0r ^ rotr_d(x0r, D) ^ S[(1 * D + r) * P] ^ S[(1 * D + dn) * P];
          S[(4 * D + r) * P] = v;
          S[(5 * D + r) * P] = pl;
          any |= v | pl;
        }
        ++attempt;
        if (any != 0u || one_draw) break;
        if (attempt >= kMaxAttempts) { st = kStatusResetGaveUp; break; }
      }
    }
  }
  __syncthreads();
  if (idx == nullptr) {                          // the CTA's envs are consecutive: coalesced copies, env-major in HBM
    const int64_t rest = n_idx - k0;
    const int cnt = rest < kResetRowsThreads ? (int)rest : kResetRowsThreads;
    uint32_t* gq = qubits + k0 * 4 * D;
    for (int g = t; g < cnt * 4 * D; g += kResetRowsThreads) gq[g] = rows_smem[(g % (4 * D)) * P + g / (4 * D)];
    uint32_t* gsyn = syn + k0 * 2 * D;
    for (int g = t; g < cnt * 2 * D; g += kResetRowsThreads)
      gsyn[g] = rows_smem[(4 * D + g % (2 * D)) * P + g / (2 * D)];
  } else if (valid) {                            // listed envs: scattered, written per thread
    uint32_t* gq = qubits + e * 4 * D;
    uint32_t* gsyn = syn + e * 2 * D;
#pragma unroll 1
    for (int w = 0; w < 4 * D; ++w) gq[w] = S[w * P];
#pragma unroll 1
    for (int w = 0; w < 2 * D; ++w) gsyn[w] = S[(4 * D + w) * P];
  }
  if (valid) {
    episode[e] = ep + 1;
    if (gs) gs[e] = 1;
    if (status && st) status[e] |= st;
  }
}

}  // namespace tc
