// libtoric_b200.so -- kernels and C ABI (include/toric_b200.h) of the toric-code environment engine.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include "../../include/toric_b200.h"

#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "fused_step.cuh"
#include "persp_write.cuh"
#include "policy_kernels.cuh"
#include "reset_env.cuh"
#include "state_kernels.cuh"

namespace tc {

// ------------------------------------------------------------------------------------------------
// mask injection / read-back
// ------------------------------------------------------------------------------------------------
// one thread per (env, matrix, row)
__global__ void pack_qubits_kernel(const uint8_t* __restrict__ codes, uint32_t* __restrict__ qubits, int d,
                                   int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 2 * d) return;
  const int64_t e = t / (2 * d);
  const int mr = (int)(t - e * 2 * d);
  const int m = mr / d, i = mr - m * d;
  const uint8_t* src = codes + t * d;
  uint32_t x = 0, z = 0;
  for (int j = 0; j < d; ++j) {
    const uint32_t c = src[j] & 3u;
    x |= (uint32_t)(c == 1u || c == 2u) << j;
    z |= (uint32_t)(c >= 2u) << j;
  }
  qubits[(e * 4 + m) * d + i] = x;
  qubits[(e * 4 + 2 + m) * d + i] = z;
}

__global__ void unpack_qubits_kernel(const uint32_t* __restrict__ qubits, uint8_t* __restrict__ codes, int d,
                                     int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 2 * d) return;
  const int64_t e = t / (2 * d);
  const int mr = (int)(t - e * 2 * d);
  const int m = mr / d, i = mr - m * d;
  const uint32_t x = qubits[(e * 4 + m) * d + i], z = qubits[(e * 4 + 2 + m) * d + i];
  uint8_t* dst = codes + t * d;
  for (int j = 0; j < d; ++j) {
    const uint32_t xb = (x >> j) & 1u, zb = (z >> j) & 1u;
    dst[j] = (uint8_t)(xb ? (zb ? 2 : 1) : (zb ? 3 : 0));
  }
}

// one thread per (env, plane, row)
__global__ void pack_state_kernel(const uint8_t* __restrict__ st, uint32_t* __restrict__ syn, int d, int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 2 * d) return;
  const uint8_t* src = st + t * d;
  uint32_t w = 0;
  for (int j = 0; j < d; ++j) w |= (uint32_t)(src[j] != 0) << j;
  syn[t] = w;
}

// 16 output bytes per thread, written as one 16-byte store: byte b of the flat [n][2][d][d] array is bit b % d of
// row word b / d.  The (at most 6, d >= 3) row words a thread needs are all loaded before the first use.
__global__ void unpack_state_kernel(const uint32_t* __restrict__ syn, uint8_t* __restrict__ st, int d, int64_t total,
                                    int64_t n_rows) {
  const int64_t b0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (b0 >= total) return;
  int64_t row;
  if (total <= 0x7fffffff) row = (int64_t)((uint32_t)b0 / (uint32_t)d);   // 32-bit divide: a 64-bit one is a subroutine
  else row = b0 / d;
  int j = (int)(b0 - row * d);
  uint32_t w[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) w[k] = row + k < n_rows ? syn[row + k] : 0u;
  // the thread's 16 stream bits (row fields of d bits, back to back), then 4 bits -> 4 bytes by a multiply
  uint32_t bits = w[0] >> j;
  int sh = d - j;
#pragma unroll
  for (int k = 1; k < 6; ++k) {
    if (sh < 16) bits |= w[k] << sh;
    sh += d;
  }
  uint32_t out[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) out[q] = (((bits >> (4 * q)) & 15u) * 0x00204081u) & 0x01010101u;
  if (b0 + 16 <= total && (((uintptr_t)st) & 15) == 0) {
    *reinterpret_cast<uint4*>(st + b0) = make_uint4(out[0], out[1], out[2], out[3]);
  } else {
    for (int k = 0; k < 16 && b0 + k < total; ++k) st[b0 + k] = (uint8_t)(out[k >> 2] >> (8 * (k & 3)));
  }
}

// row words -> np.packbits(state, axis=-1, bitorder="little"): ceil(d/8) bytes per lattice row
__global__ void state_rows_kernel(const uint32_t* __restrict__ syn, uint8_t* __restrict__ rows, int d, int nb,
                                  int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 2 * d * nb) return;
  const int64_t r = t / nb;
  const int b = (int)(t - r * nb);
  rows[t] = (uint8_t)(syn[r] >> (8 * b));
}

// ------------------------------------------------------------------------------------------------
// K2 syndrome, K3 step, K1 reset, K5 ground state (standalone forms)
// ------------------------------------------------------------------------------------------------
// The stand-alone state kernels move ~100-170 bytes per env with one row word per lane; one env per group leaves too
// few bytes in flight per SM to cover the HBM latency, so every group handles kEnvsPerGroup envs (e, e + groups,
// ...) and issues all their loads before the first use (2^22 envs, d=7, 1 -> 4 envs per group: syndrome 4.85 -> 5.93
// TB/s, ground state 3.60 -> 5.26 TB/s).
#ifndef TC_ENVS_PER_GROUP
#define TC_ENVS_PER_GROUP 4
#endif
constexpr int kEnvsPerGroup = TC_ENVS_PER_GROUP;
inline int64_t groups_for(int64_t n) { return (n + kEnvsPerGroup - 1) / kEnvsPerGroup; }
#ifndef TC_STEP_ENVS_PER_GROUP
#define TC_STEP_ENVS_PER_GROUP 2       /* the step kernel carries twice the state per env */
#endif
constexpr int kStepEnvsPerGroup = TC_STEP_ENVS_PER_GROUP;

template <int G>
__global__ void __launch_bounds__(kThreads) syndrome_kernel(const uint32_t* __restrict__ qubits,
                                                            uint32_t* __restrict__ syn, int d, int64_t n) {
  const uint32_t gmask = group_mask<G>();
  const int r = threadIdx.x & (G - 1);
  const int64_t g = ((int64_t)blockIdx.x * kThreads + threadIdx.x) / G;
  const int64_t groups = (n + kEnvsPerGroup - 1) / kEnvsPerGroup;
  Rows s[kEnvsPerGroup];
  int64_t e[kEnvsPerGroup];
  bool valid[kEnvsPerGroup];
#pragma unroll
  for (int u = 0; u < kEnvsPerGroup; ++u) {
    e[u] = g + u * groups;
    valid[u] = g < groups && e[u] < n;
    if (!valid[u]) e[u] = n - 1;
    load_qubits<G>(qubits, e[u], d, r, s[u]);
  }
#pragma unroll
  for (int u = 0; u < kEnvsPerGroup; ++u) {
    syndrome_rows<G>(gmask, d, r, s[u]);
    if (valid[u]) store_syn<G>(syn, e[u], d, r, s[u]);
  }
}

template <int G>
__global__ void __launch_bounds__(kThreads)
step_kernel(uint32_t* __restrict__ qubits, uint32_t* __restrict__ syn, const int32_t* __restrict__ actions,
            float* __restrict__ reward, uint8_t* __restrict__ terminal, uint8_t* __restrict__ status, int d,
            int64_t n) {
  const uint32_t gmask = group_mask<G>();
  const int r = threadIdx.x & (G - 1);
  const int64_t g = ((int64_t)blockIdx.x * kThreads + threadIdx.x) / G;
  const int64_t groups = (n + kStepEnvsPerGroup - 1) / kStepEnvsPerGroup;
  Rows s[kStepEnvsPerGroup];
  int64_t e[kStepEnvsPerGroup];
  bool valid[kStepEnvsPerGroup];
  int4 a[kStepEnvsPerGroup];
#pragma unroll
  for (int u = 0; u < kStepEnvsPerGroup; ++u) {
    e[u] = g + u * groups;
    valid[u] = g < groups && e[u] < n;
    if (!valid[u]) e[u] = n - 1;
    load_qubits<G>(qubits, e[u], d, r, s[u]);
    load_syn<G>(syn, e[u], d, r, s[u]);
    a[u] = *reinterpret_cast<const int4*>(actions + e[u] * 4);
  }
#pragma unroll
  for (int u = 0; u < kStepEnvsPerGroup; ++u) {
    const int before = gsum<G>(gmask, __popc(s[u].v) + __popc(s[u].pl));
    const bool ok = apply_action<G>(d, r, a[u].x, a[u].y, a[u].z, a[u].w, s[u]);
    const int after = gsum<G>(gmask, __popc(s[u].v) + __popc(s[u].pl));
    if (!valid[u]) continue;        // uniform inside the group
    store_qubits<G>(qubits, e[u], d, r, s[u]);
    store_syn<G>(syn, e[u], d, r, s[u]);
    if (r == 0) {
      // getReward / isTerminalState (ToricCode.py:202-218)
      reward[e[u]] = after == 0 ? 100.0f : (float)(before - after);
      terminal[e[u]] = after == 0 ? 1 : 0;
      if (status && !ok) status[e[u]] |= kStatusBadAction;
    }
  }
}

#ifndef TC_RESET_MIN_CTAS
#define TC_RESET_MIN_CTAS 5   /* the sampler is ALU-bound (profiles/r02_ncu_reset_d7.txt): 40 warps per SM instead of 24 */
#endif
// D is a compile-time lattice size: the sampler (inlined) then has constant slot counts, masks and call counts --
// the kernel is bound by integer instruction issue (profiles/r02_ncu_reset_d7*.txt), so this is where its time goes.
template <int G, int D>
__global__ void __launch_bounds__(kThreads, TC_RESET_MIN_CTAS)
reset_kernel(uint32_t* __restrict__ qubits, uint32_t* __restrict__ syn, uint8_t* __restrict__ gs,
             uint8_t* __restrict__ status, const uint32_t* __restrict__ p_thresh, uint32_t* __restrict__ episode,
             const int64_t* __restrict__ idx, int64_t n_idx, int mode, int n_fixed, uint2 key, int64_t env_base) {
  constexpr int d = D;
  const uint32_t gmask = group_mask<G>();
  const int r = threadIdx.x & (G - 1);
  const int64_t k = ((int64_t)blockIdx.x * kThreads + threadIdx.x) / G;
  if (k >= n_idx) return;   // whole groups leave together
  const int64_t e = idx ? idx[k] : k;
  const uint32_t ep = episode[e];
  Rows s;
  uint8_t st = 0;
  SampleStats stats;
  sample_errors<G>(gmask, d, r, (uint64_t)(env_base + e), ep, p_thresh[e], mode, n_fixed, key, s, st, stats);
  store_qubits<G>(qubits, e, d, r, s);
  store_syn<G>(syn, e, d, r, s);
  __syncwarp(gmask);
  if (r == 0) {
    episode[e] = ep + 1;
    if (gs) gs[e] = 1;
    if (status && st) status[e] |= st;
  }
}

template <int G>
__global__ void __launch_bounds__(kThreads) ground_state_kernel(const uint32_t* __restrict__ qubits,
                                                                uint8_t* __restrict__ gs, int d, int64_t n) {
  const uint32_t gmask = group_mask<G>();
  const int r = threadIdx.x & (G - 1);
  const int64_t g = ((int64_t)blockIdx.x * kThreads + threadIdx.x) / G;
  const int64_t groups = (n + kEnvsPerGroup - 1) / kEnvsPerGroup;
  Rows s[kEnvsPerGroup];
  int64_t e[kEnvsPerGroup];
  bool valid[kEnvsPerGroup];
#pragma unroll
  for (int u = 0; u < kEnvsPerGroup; ++u) {
    e[u] = g + u * groups;
    valid[u] = g < groups && e[u] < n;
    if (!valid[u]) e[u] = n - 1;
    load_qubits<G>(qubits, e[u], d, r, s[u]);
  }
#pragma unroll
  for (int u = 0; u < kEnvsPerGroup; ++u) {
    const bool even = parity_even<G>(gmask, s[u]);
    if (valid[u] && r == 0 && !even) gs[e[u]] = 0;   // sticky (ToricCode.py:332-335)
  }
}

// ------------------------------------------------------------------------------------------------
// perspective count + CTA-local exclusive prefix  (first half of K4)
// ------------------------------------------------------------------------------------------------
// CTA size of the kernels that produce the CTA-local perspective prefix (actor_step_kernel, persp_count_kernel); the
// writer's PerspCfg::kEpb follows it.  Every CTA ends in a barrier, and a CTA that holds an auto-reset keeps the others
// of its SM slot waiting, so smaller CTAs finish more evenly; the scan kernel then has proportionally more sums to add.
#ifndef TC_COUNT_THREADS
#define TC_COUNT_THREADS 256
#endif
constexpr int kCountThreads = TC_COUNT_THREADS;
// cta_count_prefix scans the CTA's env counts in ONE warp out of `sc[32]`: at most 32 envs per CTA at the narrowest
// group (G = 8); whole warps only; and actor_step_kernel's launch bound scales by 256 / kCountThreads
static_assert(kCountThreads % 32 == 0 && kCountThreads >= 32, "TC_COUNT_THREADS: whole warps");
static_assert(kCountThreads / 8 <= 32, "TC_COUNT_THREADS: more than 32 envs per CTA overruns cta_count_prefix's scan");
static_assert(256 % kCountThreads == 0, "TC_COUNT_THREADS must divide 256 (launch bounds of actor_step_kernel)");

template <int G>
__device__ __forceinline__ void cta_count_prefix(int count, bool valid, int r, int64_t e,
                                                 int32_t* __restrict__ counts, int32_t* __restrict__ local_off,
                                                 int32_t* __restrict__ block_sum) {
  constexpr int EPB = kCountThreads / G;
  __shared__ int sc[32];
  const int ge = threadIdx.x / G;
  if (r == 0) sc[ge] = valid ? count : 0;
  __syncthreads();
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const int v = lane < EPB ? sc[lane] : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane < EPB) sc[lane] = inc - v;
    if (lane == 31) block_sum[blockIdx.x] = inc;
  }
  __syncthreads();
  if (valid && r == 0) {
    counts[e] = count;
    local_off[e] = sc[ge];
  }
}

template <int G>
__global__ void __launch_bounds__(kCountThreads)
persp_count_kernel(const uint32_t* __restrict__ syn, int32_t* __restrict__ counts, int32_t* __restrict__ local_off,
                   int32_t* __restrict__ block_sum, int d, int64_t n) {
  const uint32_t gmask = group_mask<G>();
  const int r = threadIdx.x & (G - 1);
  int64_t e = ((int64_t)blockIdx.x * kCountThreads + threadIdx.x) / G;
  const bool valid = e < n;
  if (!valid) e = n - 1;
  Rows s;
  load_syn<G>(syn, e, d, r, s);
  uint32_t m0, m1;
  persp_masks<G>(gmask, d, r, s.v, s.pl, m0, m1);
  const int count = gsum<G>(gmask, __popc(m0) + __popc(m1));
  cta_count_prefix<G>(count, valid, r, e, counts, local_off, block_sum);
}

// Exclusive scan of the per-CTA sums -> block_off (int64), grand total -> offsets[n].
// One CTA handles up to kScanChunk sums: warp w owns a contiguous piece of them and walks it in 32-wide tiles (coalesced
// 128-byte reads, no CTA barrier inside the loops): pass 1 adds the piece up, the 32 piece totals are scanned once,
// pass 2 re-reads the piece and writes the running prefix.  Batches with more than kScanChunk count-CTAs (2^18 envs at
// d <= 8) first reduce every chunk (scan_chunk_sums_kernel) and then run one scan CTA per chunk, each starting from the
// sum of the chunks in front of it -- a single CTA walking 131 072 sums was 0.25 ms at 2^22 envs.
constexpr int kScanChunk = 8192;

__global__ void __launch_bounds__(1024)
scan_chunk_sums_kernel(const int32_t* __restrict__ block_sum, int64_t nblocks, long long* __restrict__ chunk_sum) {
  __shared__ long long warp_tot[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_wait();
  pdl_launch_dependents();
  const int64_t c0 = (int64_t)blockIdx.x * kScanChunk;
  const int64_t c1 = c0 + kScanChunk < nblocks ? c0 + kScanChunk : nblocks;
  long long sum = 0;
  for (int64_t i = c0 + tid; i < c1; i += 1024) sum += (long long)block_sum[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) warp_tot[warp] = sum;
  __syncthreads();
  if (warp == 0) {
    sum = warp_tot[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) chunk_sum[blockIdx.x] = sum;
  }
}

__global__ void __launch_bounds__(1024)
scan_blocks_kernel(const int32_t* __restrict__ block_sum, int64_t* __restrict__ block_off, int64_t nblocks,
                   int64_t* __restrict__ total_out, int64_t* __restrict__ counters,
                   const long long* __restrict__ chunk_sum) {
  __shared__ long long warp_tot[32];
  __shared__ long long s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_wait();                 // block_sum (and chunk_sum) come from the kernels in front
  pdl_launch_dependents();
  // this CTA's sums [c0, c1); with a single CTA that is everything (chunk_sum == nullptr)
  const int64_t c0 = chunk_sum ? (int64_t)blockIdx.x * kScanChunk : 0;
  const int64_t c1 = chunk_sum ? (c0 + kScanChunk < nblocks ? c0 + kScanChunk : nblocks) : nblocks;
  if (warp == 0) {            // sum of the chunks in front (at most a few dozen)
    long long b = 0;
    if (chunk_sum)
      for (int c = lane; c < (int)blockIdx.x; c += 32) b += chunk_sum[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) b += __shfl_xor_sync(0xffffffffu, b, o);
    if (lane == 0) s_base = b;
  }
  const int64_t len = c1 - c0;
  const int64_t piece = ((len + 1023) / 1024) * 32;              // multiple of the tile width
  const int64_t lo = c0 + ((int64_t)warp * piece < len ? (int64_t)warp * piece : len);
  const int64_t hi = lo + piece < c1 ? lo + piece : c1;
  long long sum = 0;
#pragma unroll 4
  for (int64_t i = lo + lane; i < hi; i += 32) sum += (long long)block_sum[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) warp_tot[warp] = sum;
  __syncthreads();
  const long long base = s_base;
  if (warp == 0) {
    const long long w = warp_tot[lane];
    long long winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    warp_tot[lane] = winc - w;   // exclusive prefix of the piece totals
    if (lane == 31 && c1 == nblocks) {          // the CTA that holds the last sums knows the grand total
      *total_out = base + winc;
      if (counters) atomicAdd(reinterpret_cast<unsigned long long*>(counters + kCtrPerspectives),
                              (unsigned long long)(base + winc));
    }
  }
  __syncthreads();
  long long carry = base + warp_tot[warp];
  for (int64_t t0 = lo; t0 < hi; t0 += 32) {
    const int64_t i = t0 + lane;
    const int v = i < hi ? block_sum[i] : 0;
    int inc = v;                 // a tile of 32 CTA sums stays far below 2^31 (<= 32 * 32 * 2 * 32 * 32)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (i < hi) block_off[i] = carry + (long long)(inc - v);
    carry += (long long)__shfl_sync(0xffffffffu, inc, 31);
  }
}

// ------------------------------------------------------------------------------------------------
// fused actor iteration: [random action] -> step -> reward/terminal -> bookkeeping / auto-reset -> count
// (loop body of src/Actor_mp.py:104-185 without the network; perspectives follow in persp_write_kernel)
// ------------------------------------------------------------------------------------------------
// ActorArgs: fused_step.cuh

#ifndef TC_ACTOR_MIN_CTAS
#define TC_ACTOR_MIN_CTAS 5   /* 48 registers: 5 CTAs per SM, 2048 CTAs at 65536 envs = 2.8 waves instead of 3.5 */
#endif
template <int G>
__global__ void __launch_bounds__(kCountThreads, TC_ACTOR_MIN_CTAS * (256 / kCountThreads)) actor_step_kernel(const ActorArgs a) {
  __shared__ unsigned int s_ctr[kNumCounters];
  if (threadIdx.x < kNumCounters) s_ctr[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t gmask = group_mask<G>();
  const int d = a.d;
  const int r = threadIdx.x & (G - 1);
  int64_t e = ((int64_t)blockIdx.x * kCountThreads + threadIdx.x) / G;
  const bool valid = e < a.n;
  if (!valid) e = a.n - 1;
  const bool lead = valid && r == 0;
  pdl_wait();                 // the previous step's perspective writer still reads syn
  pdl_launch_dependents();
  Rows s;
  load_qubits<G>(a.qubits, e, d, r, s);
  load_syn<G>(a.syn, e, d, r, s);
  uint32_t ep = a.episode[e];
  int stp = a.steps[e];
  const uint64_t env_id = (uint64_t)(a.env_base + e);

  int am, ai, aj, aop;
  if (a.flags & TC_ACTOR_RANDOM_ACTIONS) {
    uint32_t m0, m1;
    persp_masks<G>(gmask, d, r, s.v, s.pl, m0, m1);
    random_action<G>(gmask, d, r, m0, m1, env_id, ep, (uint32_t)stp, a.key, am, ai, aj, aop);
    if (lead) *reinterpret_cast<int4*>(a.actions + e * 4) = make_int4(am, ai, aj, aop);
  } else {
    const int4 act = *reinterpret_cast<const int4*>(a.actions + e * 4);
    am = act.x; ai = act.y; aj = act.z; aop = act.w;
  }
  const int before = gsum<G>(gmask, __popc(s.v) + __popc(s.pl));
  // TC_ACTOR_FREEZE_DONE: an env that is already solved is left exactly as it is (batched evaluation keeps
  // finished episodes in the batch; the reference's per-episode while loop simply ends, evaluation.py:79)
  const bool frozen = (a.flags & TC_ACTOR_FREEZE_DONE) && before == 0;
  const bool ok = frozen ? true : apply_action<G>(d, r, am, ai, aj, aop, s);
  const int after = gsum<G>(gmask, __popc(s.v) + __popc(s.pl));
  const bool term = after == 0;
  const bool even = parity_even<G>(gmask, s);
  if (!frozen) stp += 1;
  if (lead && !frozen) {
    a.reward[e] = term ? 100.0f : (float)(before - after);
    a.terminal[e] = term ? 1 : 0;
    a.gs[e] = even ? 1 : 0;
    atomicAdd(&s_ctr[kCtrEnvSteps], 1u);
    if (!ok) {
      if (a.status) a.status[e] |= kStatusBadAction;
      atomicAdd(&s_ctr[kCtrBadActions], 1u);
    }
  }
  if (a.prev_syn && valid) store_syn<G>(a.prev_syn, e, d, r, s);

  const bool done = !frozen && (term || stp > a.max_steps);      // Actor_mp.py:171-172
  if ((a.flags & TC_ACTOR_AUTO_RESET) && done) {     // uniform inside the group
    uint8_t st = 0;
    SampleStats stats;
    sample_errors<G>(gmask, d, r, env_id, ep, a.p_thresh[e], a.mode, a.n_fixed, a.key, s, st, stats);
    const int nx = gsum<G>(gmask, stats.nx), ny = gsum<G>(gmask, stats.ny), nz = gsum<G>(gmask, stats.nz);
    ep += 1;
    stp = 0;
    if (lead) {
      atomicAdd(&s_ctr[kCtrEpisodes], 1u);
      if (term) atomicAdd(&s_ctr[kCtrSolved], 1u); else atomicAdd(&s_ctr[kCtrTimeouts], 1u);
      if (term && even) atomicAdd(&s_ctr[kCtrGroundKept], 1u);
      atomicAdd(&s_ctr[kCtrErrX], (unsigned)nx);
      atomicAdd(&s_ctr[kCtrErrY], (unsigned)ny);
      atomicAdd(&s_ctr[kCtrErrZ], (unsigned)nz);
      atomicAdd(&s_ctr[kCtrQubitsSampled], (unsigned)(2 * d * d));
      atomicAdd(&s_ctr[kCtrResetAttempts], (unsigned)stats.attempts);
      if (st && a.status) a.status[e] |= st;
    }
  }
  if (valid) {
    store_qubits<G>(a.qubits, e, d, r, s);
    store_syn<G>(a.syn, e, d, r, s);
    if (a.state_rows && r < d) {          // host-buffer step: the download format comes straight from the step kernel
      const int nb = (d + 7) >> 3;
      uint8_t* rv = a.state_rows + (e * 2 * d + r) * nb;
      uint8_t* rp = rv + (int64_t)d * nb;
      for (int b = 0; b < nb; ++b) {
        rv[b] = (uint8_t)(s.v >> (8 * b));
        rp[b] = (uint8_t)(s.pl >> (8 * b));
      }
    }
  }
  if (lead) {
    a.steps[e] = stp;
    a.episode[e] = ep;
  }
  uint32_t m0, m1;
  persp_masks<G>(gmask, d, r, s.v, s.pl, m0, m1);
  const int count = gsum<G>(gmask, __popc(m0) + __popc(m1));
  cta_count_prefix<G>(count, valid, r, e, a.counts, a.local_off, a.block_sum);   // has __syncthreads
  if (a.counters && threadIdx.x < kNumCounters && s_ctr[threadIdx.x])
    atomicAdd(reinterpret_cast<unsigned long long*>(a.counters + threadIdx.x), (unsigned long long)s_ctr[threadIdx.x]);
}

}  // namespace tc

// ================================================================================================
// C ABI
// ================================================================================================
namespace {

using namespace tc;

inline bool bad_d(int d) { return d < kMinD || d > kMaxD; }
inline int launch_status() {
  const cudaError_t err = cudaGetLastError();
  return err == cudaSuccess ? 0 : (int)err;
}
inline unsigned grid_for(int64_t threads, int per_block) { return (unsigned)((threads + per_block - 1) / per_block); }
// launch with the programmatic-stream-serialization attribute (see common.cuh: pdl_wait / pdl_launch_dependents)
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
                       Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(block, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = TC_PDL ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline uint2 make_key(uint64_t seed) { return make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)); }

#define TC_DISPATCH_G(d, CALL)                 \
  do {                                         \
    if ((d) <= 8) { constexpr int G = 8; CALL; }        \
    else if ((d) <= 16) { constexpr int G = 16; CALL; } \
    else { constexpr int G = 32; CALL; }                \
  } while (0)

struct Workspace {
  int32_t* local_off;
  int32_t* block_sum;
  int64_t* block_off;
  int64_t nblocks;
  int epb;
  ScanWs scan;        // single-pass step: launch counter + look-back descriptors (never touched by the other kernels)
  long long* chunk_sum;   // [ceil(nblocks / kScanChunk)]: multi-CTA scan of large batches
  int64_t nchunks;
  size_t bytes;
};

inline Workspace carve(void* ws, int d, int64_t n) {
  Workspace w;
  w.epb = kCountThreads / group_size_for(d);
  w.nblocks = (n + w.epb - 1) / w.epb;
  size_t o = 0;
  char* base = static_cast<char*>(ws);
  w.local_off = reinterpret_cast<int32_t*>(base + o);
  o += (size_t)n * 4;
  w.block_sum = reinterpret_cast<int32_t*>(base + o);
  o += (size_t)w.nblocks * 4;
  o = (o + 15) & ~(size_t)15;
  w.block_off = reinterpret_cast<int64_t*>(base + o);
  o += (size_t)w.nblocks * 8;
  o = (o + 127) & ~(size_t)127;
  w.scan.header = reinterpret_cast<uint32_t*>(base + o);
  o += 64;
  w.scan.desc1 = reinterpret_cast<unsigned long long*>(base + o);
  o += (size_t)((n + 31) / 32) * 8;
  w.scan.desc0 = reinterpret_cast<uint32_t*>(base + o);
  o += (size_t)n * 4;
  o = (o + 15) & ~(size_t)15;
  w.nchunks = (w.nblocks + kScanChunk - 1) / kScanChunk;
  w.chunk_sum = reinterpret_cast<long long*>(base + o);
  o += (size_t)w.nchunks * 8;
  w.bytes = (o + 255) & ~(size_t)255;
  return w;
}

// cross-CTA scan of the perspective counts: one CTA up to kScanChunk count-CTAs, else chunk sums + one CTA per chunk
inline void launch_scan(const Workspace& w, int64_t* total_out, int64_t* counters, cudaStream_t st, bool pdl) {
  if (w.nblocks <= kScanChunk) {
    if (pdl)
      launch_pdl(scan_blocks_kernel, 1, 1024, 0, st, (const int32_t*)w.block_sum, w.block_off, w.nblocks, total_out,
                 counters, (const long long*)nullptr);
    else
      scan_blocks_kernel<<<1, 1024, 0, st>>>(w.block_sum, w.block_off, w.nblocks, total_out, counters, nullptr);
    return;
  }
  if (pdl) {
    launch_pdl(scan_chunk_sums_kernel, (unsigned)w.nchunks, 1024, 0, st, (const int32_t*)w.block_sum, w.nblocks,
               w.chunk_sum);
    launch_pdl(scan_blocks_kernel, (unsigned)w.nchunks, 1024, 0, st, (const int32_t*)w.block_sum, w.block_off,
               w.nblocks, total_out, counters, (const long long*)w.chunk_sum);
  } else {
    scan_chunk_sums_kernel<<<(unsigned)w.nchunks, 1024, 0, st>>>(w.block_sum, w.nblocks, w.chunk_sum);
    scan_blocks_kernel<<<(unsigned)w.nchunks, 1024, 0, st>>>(w.block_sum, w.block_off, w.nblocks, total_out, counters,
                                                            w.chunk_sum);
  }
}

// persistent grid: one CTA slot per (SM x resident CTAs), capped by the number of envs
template <int D, typename OutT>
int launch_write(const uint32_t* syn, const Workspace& w, OutT* out, uint8_t* pos, int64_t* offsets,
                  int32_t* overflow, int64_t cap, int64_t n, uint32_t one16, cudaStream_t st) {
  using Cfg = PerspCfg<D>;
  constexpr int WPB = Cfg::kWarps;
  // per device and instantiation, once (thread-safe): opt in to the dynamic shared memory, SMs x resident CTAs
  static int slots[64] = {0};
  static std::once_flag once[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  std::call_once(once[dev], [dev] {
    cudaFuncSetAttribute(persp_write_kernel<D, OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         Cfg::kSmemBytes);
    int sms = 148, occ = 1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, persp_write_kernel<D, OutT>, WPB * 32, Cfg::kSmemBytes);
    slots[dev] = sms * (occ > 0 ? occ : 1);
  });
  // TC_WRITE_WAVES: CTAs per resident slot.  0 = one env per warp (the hardware CTA scheduler staggers the
  // env setups against the streaming of the others); k > 0 = persistent warps walking ~n/(slots*k*WPB) envs.
#ifndef TC_WRITE_WAVES
#define TC_WRITE_WAVES 0
#endif
  // TC_WRITE_SPLIT: warps per env on the row path (d > 8), where one env is tens to hundreds of 4 KB tiles:
  // enough (env, slice) items for >= 8 waves of CTAs, at most 16 slices and never more slices than an env of
  // average density has tiles.  The plane path (d <= 8: ~6 tiles per env, setup-heavy) keeps one warp per env.
#ifndef TC_WRITE_SPLIT
#define TC_WRITE_SPLIT 0          /* 0 = automatic */
#endif
  if (w.epb != Cfg::kEpb) return -1;          // the count kernels' CTA shape is baked into the writer
#ifndef TC_NO_GROUP_WRITER
  if constexpr (Cfg::kGroup > 1) {            // small lattices: kGroup consecutive envs per warp, one stream
    static std::once_flag once_g[64];
    std::call_once(once_g[dev], [] {
      cudaFuncSetAttribute(persp_write_group_kernel<D, OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           Cfg::kSmemBytes);
    });
    const int64_t items = (n + Cfg::kGroup - 1) / Cfg::kGroup;
    launch_pdl(persp_write_group_kernel<D, OutT>, grid_for(items, WPB), WPB * 32, Cfg::kSmemBytes, st, syn,
               (const int32_t*)w.local_off, (const int64_t*)w.block_off, out, pos, offsets, overflow, cap, n, one16);
    return 0;
  }
#endif
  int split = 1;
  if (!Cfg::kPlane) {
    if (TC_WRITE_SPLIT > 0) {
      split = TC_WRITE_SPLIT;
    } else {
      const int64_t ctas = (n + WPB - 1) / WPB;
      const int64_t want = ((int64_t)slots[dev] * 8 + ctas - 1) / ctas;
      const int tiles = (D * D / 2) * 2 * D * D * (int)sizeof(OutT) / kTileBytes;   // env with P = d^2/2
      split = (int)(want < 1 ? 1 : (want > 16 ? 16 : want));
      if (split > tiles) split = tiles > 1 ? tiles : 1;
    }
  }
  unsigned grid = grid_for(n * split, WPB);
  if (TC_WRITE_WAVES > 0 && grid > (unsigned)slots[dev] * TC_WRITE_WAVES) grid = (unsigned)slots[dev] * TC_WRITE_WAVES;
  launch_pdl(persp_write_kernel<D, OutT>, grid, WPB * 32, Cfg::kSmemBytes, st, syn, (const int32_t*)w.local_off,
             (const int64_t*)w.block_off, w.epb, out, pos, offsets, overflow, cap, n, one16, split);
  return 0;
}

template <typename OutT>
int write_dispatch(int d, const uint32_t* syn, const Workspace& w, OutT* out, uint8_t* pos, int64_t* offsets,
                   int32_t* overflow, int64_t cap, int64_t n, uint32_t one16, cudaStream_t st) {
  switch (d) {
#define TC_CASE(D) \
  case D:          \
    if (launch_write<D, OutT>(syn, w, out, pos, offsets, overflow, cap, n, one16, st) != 0) return -1; \
    break;
    TC_CASE(3) TC_CASE(4) TC_CASE(5) TC_CASE(6) TC_CASE(7) TC_CASE(8) TC_CASE(9) TC_CASE(10)
    TC_CASE(11) TC_CASE(12) TC_CASE(13) TC_CASE(14) TC_CASE(15) TC_CASE(16) TC_CASE(17) TC_CASE(18)
    TC_CASE(19) TC_CASE(20) TC_CASE(21) TC_CASE(22) TC_CASE(23) TC_CASE(24) TC_CASE(25) TC_CASE(26)
    TC_CASE(27) TC_CASE(28) TC_CASE(29) TC_CASE(30) TC_CASE(31) TC_CASE(32)
#undef TC_CASE
    default: return -1;
  }
  return launch_status();
}

// single-pass step (fused_step.cuh): one env per warp, CTAs in env order
template <int D>
int launch_fused(const ActorArgs& a, const Workspace& w, float* out, uint8_t* pos, int64_t* offsets,
                 int32_t* overflow, int64_t cap, cudaStream_t st) {
  using Cfg = PerspCfg<D>;
  static std::once_flag once[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  std::call_once(once[dev], [] {
    cudaFuncSetAttribute(fused_step_kernel<D, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
  });
  launch_pdl(fused_step_kernel<D, float>, grid_for(a.n, Cfg::kWarps), Cfg::kWarps * 32, Cfg::kSmemBytes, st, a,
             w.scan, out, pos, offsets, overflow, cap, 0u);
  return 0;
}

int fused_dispatch(const ActorArgs& a, const Workspace& w, float* out, uint8_t* pos, int64_t* offsets,
                   int32_t* overflow, int64_t cap, cudaStream_t st) {
  switch (a.d) {
#define TC_CASE(D) \
  case D:          \
    launch_fused<D>(a, w, out, pos, offsets, overflow, cap, st); \
    break;
    TC_CASE(3) TC_CASE(4) TC_CASE(5) TC_CASE(6) TC_CASE(7) TC_CASE(8) TC_CASE(9) TC_CASE(10)
    TC_CASE(11) TC_CASE(12) TC_CASE(13) TC_CASE(14) TC_CASE(15) TC_CASE(16)
#undef TC_CASE
    default: return -1;
  }
  return launch_status();
}

// tc_actor_pre with the optional row-bitmask output of the host-buffer step
int actor_pre_impl(uint32_t* qubits, uint32_t* syn, uint32_t* prev_syn, int32_t* actions, float* reward,
                 uint8_t* terminal, uint8_t* gs, uint8_t* status, int32_t* steps, uint32_t* episode,
                 const uint32_t* p_thresh, int64_t* counters, int flags, int max_steps, int mode, int n_fixed,
                 uint64_t seed, int64_t env_base, int32_t* counts, int64_t* offsets, void* workspace,
                 size_t workspace_bytes, int d, int64_t n, void* stream, uint8_t* state_rows) {
  if (bad_d(d) || n < 0 || !offsets) return -1;
  if (mode < 0 || mode > 2 || n_fixed < 0 || (mode != 0 && n_fixed == 0)) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    cudaMemsetAsync(offsets, 0, sizeof(int64_t), st);
    return launch_status();
  }
  if (!qubits || !syn || !actions || !reward || !terminal || !gs || !steps || !episode || !p_thresh || !counts ||
      !workspace)
    return -1;
  if ((uintptr_t)actions & 15) return -1;
  const Workspace w = carve(workspace, d, n);
  if (workspace_bytes < w.bytes) return -2;
  ActorArgs a;
  a.qubits = qubits; a.syn = syn; a.prev_syn = prev_syn; a.actions = actions; a.reward = reward;
  a.terminal = terminal; a.gs = gs; a.status = status; a.steps = steps; a.episode = episode;
  a.p_thresh = p_thresh; a.counters = counters; a.counts = counts; a.local_off = w.local_off;
  a.block_sum = w.block_sum; a.key = make_key(seed); a.env_base = env_base; a.n = n; a.flags = flags;
  a.max_steps = max_steps; a.mode = mode; a.n_fixed = n_fixed; a.d = d; a.state_rows = state_rows;
  TC_DISPATCH_G(d, (launch_pdl(actor_step_kernel<G>, (unsigned)w.nblocks, kCountThreads, 0, st, a)));
  launch_scan(w, offsets + n, counters, st, true);
  return launch_status();
}

// ---- built-in scratch of tc_perspectives_static -----------------------------------------------------------------
// Lives in the library image: every GPU that loads libtoric_b200.so holds TC_STATIC_SCRATCH_BYTES of it until the
// process exits, whether or not the entry point is ever called.  One array per device, shared by all callers.
static __device__ __align__(256) unsigned char g_static_scratch[TC_STATIC_SCRATCH_BYTES];

// Device address of the CURRENT device's copy of g_static_scratch.  On the host the name `g_static_scratch` is only
// a handle for the runtime's symbol table -- handing it to a kernel would pass a host address -- so the real address
// is asked for once per device (with that device current) and remembered.  A failed lookup is not remembered: the
// next call tries again.  Devices past the table are looked up on every call.
int static_scratch(void** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static void* addr[kMaxDev] = {nullptr};
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  const bool cached = dev >= 0 && dev < kMaxDev;
  std::lock_guard<std::mutex> lock(mu);
  if (cached && addr[dev]) {
    *out = addr[dev];
    return 0;
  }
  void* p = nullptr;
  err = cudaGetSymbolAddress(&p, g_static_scratch);
  if (err != cudaSuccess) return (int)err;
  if (!p) return (int)cudaErrorInvalidSymbol;
  if (cached) addr[dev] = p;
  *out = p;
  return 0;
}

}  // namespace

extern "C" {

int tc_abi_version(void) { return TC_ABI_VERSION; }

int tc_workspace_bytes(int d, int64_t n, size_t* out_bytes) {
  if (bad_d(d) || n < 0 || !out_bytes) return -1;
  *out_bytes = carve(nullptr, d, n).bytes;
  return 0;
}

int tc_pack_qubits(const uint8_t* codes, uint32_t* qubits, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!codes || !qubits) return -1;
  pack_qubits_kernel<<<grid_for(n * 2 * d, 256), 256, 0, (cudaStream_t)stream>>>(codes, qubits, d, n);
  return launch_status();
}

int tc_unpack_qubits(const uint32_t* qubits, uint8_t* codes, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!codes || !qubits) return -1;
  unpack_qubits_kernel<<<grid_for(n * 2 * d, 256), 256, 0, (cudaStream_t)stream>>>(qubits, codes, d, n);
  return launch_status();
}

int tc_pack_state(const uint8_t* state01, uint32_t* syn, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!state01 || !syn) return -1;
  pack_state_kernel<<<grid_for(n * 2 * d, 256), 256, 0, (cudaStream_t)stream>>>(state01, syn, d, n);
  return launch_status();
}

int tc_unpack_state(const uint32_t* syn, uint8_t* state01, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!state01 || !syn) return -1;
  const int64_t total = n * 2 * d * d;
  unpack_state_kernel<<<grid_for((total + 15) / 16, 256), 256, 0, (cudaStream_t)stream>>>(syn, state01, d, total,
                                                                                        n * 2 * d);
  return launch_status();
}

int tc_state_rows(const uint32_t* syn, uint8_t* rows, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!rows || !syn) return -1;
  const int nb = (d + 7) / 8;
  state_rows_kernel<<<grid_for(n * 2 * d * nb, 256), 256, 0, (cudaStream_t)stream>>>(syn, rows, d, nb, n);
  return launch_status();
}

int tc_syndrome(const uint32_t* qubits, uint32_t* syn, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!qubits || !syn) return -1;
  TC_DISPATCH_G(d, (syndrome_kernel<G><<<grid_for(groups_for(n) * G, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                       qubits, syn, d, n)));
  return launch_status();
}

int tc_step(uint32_t* qubits, uint32_t* syn, const int32_t* actions, float* reward, uint8_t* terminal,
            uint8_t* status, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;
  if (!qubits || !syn || !actions || !reward || !terminal) return -1;
  if ((uintptr_t)actions & 15) return -1;
  const int64_t step_groups = (n + kStepEnvsPerGroup - 1) / kStepEnvsPerGroup;
  TC_DISPATCH_G(d, (step_kernel<G><<<grid_for(step_groups * G, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                       qubits, syn, actions, reward, terminal, status, d, n)));
  return launch_status();
}

int tc_reset(uint32_t* qubits, uint32_t* syn, uint8_t* gs, uint8_t* status, const uint32_t* p_thresh,
             uint32_t* episode, const int64_t* idx, int64_t n_idx, int mode, int n_fixed, uint64_t seed,
             int64_t env_base, int d, void* stream) {
  const int base_mode = mode & ~TC_SAMPLE_ONE_DRAW;
  if (bad_d(d) || n_idx < 0) return -1;
  if (base_mode < 0 || base_mode > 2 || n_fixed < 0 || (base_mode != 0 && n_fixed == 0)) return -1;
  if (n_idx == 0) return 0;
  if (!qubits || !syn || !p_thresh || !episode) return -1;
#ifndef TC_NO_RESET_ENV_KERNEL
  if (base_mode == 0 && d <= 8) {     // small lattices, depolarizing: one thread per env (csrc/reset_env.cuh)
    const int one_draw = (mode & TC_SAMPLE_ONE_DRAW) ? 1 : 0;
    const unsigned grid = grid_for(n_idx, kResetEnvThreads);
    cudaStream_t st = (cudaStream_t)stream;
    switch (d) {
#define TC_CASE(D)                                                                                                    \
  case D:                                                                                                             \
    reset_env_kernel<D><<<grid, kResetEnvThreads, 0, st>>>(qubits, syn, gs, status, p_thresh, episode, idx, n_idx,   \
                                                            one_draw, make_key(seed), env_base);                      \
    break;
      TC_CASE(3) TC_CASE(4) TC_CASE(5) TC_CASE(6) TC_CASE(7) TC_CASE(8)
#undef TC_CASE
    }
    return launch_status();
  }
  if (base_mode == 0 && d <= 16) {    // 9..16: one thread per env, rows in shared memory (csrc/reset_env.cuh)
    const int one_draw = (mode & TC_SAMPLE_ONE_DRAW) ? 1 : 0;
    const unsigned grid = grid_for(n_idx, kResetRowsThreads);
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    switch (d) {
#define TC_CASE(D)                                                                                                      \
  case D: {                                                                                                             \
    static std::once_flag once[64];                                                                                     \
    std::call_once(once[dev], [] {                                                                                      \
      cudaFuncSetAttribute(reset_rows_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize,                           \
                           reset_rows_smem_bytes<D>());                                                                 \
    });                                                                                                                 \
    reset_rows_kernel<D><<<grid, kResetRowsThreads, reset_rows_smem_bytes<D>(), st>>>(                                  \
        qubits, syn, gs, status, p_thresh, episode, idx, n_idx, one_draw, make_key(seed), env_base);                    \
  } break;
      TC_CASE(9) TC_CASE(10) TC_CASE(11) TC_CASE(12) TC_CASE(13) TC_CASE(14) TC_CASE(15) TC_CASE(16)
#undef TC_CASE
    }
    return launch_status();
  }
#endif
  switch (d) {
#define TC_CASE(D)                                                                                              \
  case D: {                                                                                                     \
    constexpr int G = D <= 8 ? 8 : (D <= 16 ? 16 : 32);                                                         \
    reset_kernel<G, D><<<grid_for(n_idx * G, kThreads), kThreads, 0, (cudaStream_t)stream>>>(                   \
        qubits, syn, gs, status, p_thresh, episode, idx, n_idx, mode, n_fixed, make_key(seed), env_base);       \
  } break;
    TC_CASE(3) TC_CASE(4) TC_CASE(5) TC_CASE(6) TC_CASE(7) TC_CASE(8) TC_CASE(9) TC_CASE(10)
    TC_CASE(11) TC_CASE(12) TC_CASE(13) TC_CASE(14) TC_CASE(15) TC_CASE(16) TC_CASE(17) TC_CASE(18)
    TC_CASE(19) TC_CASE(20) TC_CASE(21) TC_CASE(22) TC_CASE(23) TC_CASE(24) TC_CASE(25) TC_CASE(26)
    TC_CASE(27) TC_CASE(28) TC_CASE(29) TC_CASE(30) TC_CASE(31) TC_CASE(32)
#undef TC_CASE
    default: return -1;
  }
  return launch_status();
}

int tc_ground_state(const uint32_t* qubits, uint8_t* gs, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (n == 0) return 0;            // an empty batch is legal (and has no buffers)
  if (!qubits || !gs) return -1;
  TC_DISPATCH_G(d, (ground_state_kernel<G><<<grid_for(groups_for(n) * G, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                       qubits, gs, d, n)));
  return launch_status();
}

int tc_persp_count(const uint32_t* syn, int32_t* counts, int64_t* offsets, void* workspace,
                   size_t workspace_bytes, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0 || !offsets) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {               // empty batch: offsets = [0], nothing else is touched
    cudaMemsetAsync(offsets, 0, sizeof(int64_t), st);
    return launch_status();
  }
  if (!syn || !counts || !workspace) return -1;
  const Workspace w = carve(workspace, d, n);
  if (workspace_bytes < w.bytes) return -2;
  TC_DISPATCH_G(d, (persp_count_kernel<G><<<(unsigned)w.nblocks, kCountThreads, 0, st>>>(syn, counts, w.local_off,
                                                                                    w.block_sum, d, n)));
  launch_scan(w, offsets + n, nullptr, st, false);
  return launch_status();
}

int tc_persp_write_as(const uint32_t* syn, void* out, int out_dtype, uint8_t* pos, int64_t* offsets,
                      int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n,
                      void* stream) {
  if (bad_d(d) || n < 0 || cap < 0) return -1;
  if (out_dtype < TC_DTYPE_F32 || out_dtype > TC_DTYPE_F16) return -1;
  if (n == 0) return 0;
  if (!syn || !offsets || !workspace) return -1;
  if (cap > 0 && !out) return -1;
  if (((uintptr_t)out & 15) || ((uintptr_t)pos & 3)) return -1;
  const Workspace w = carve(workspace, d, n);
  if (workspace_bytes < w.bytes) return -2;
  if (n == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  switch (out_dtype) {
    case TC_DTYPE_F32: return write_dispatch<float>(d, syn, w, (float*)out, pos, offsets, overflow, cap, n, 0u, st);
    case TC_DTYPE_U8:
      return write_dispatch<uint8_t>(d, syn, w, (uint8_t*)out, pos, offsets, overflow, cap, n, 0u, st);
    case TC_DTYPE_BF16:
      return write_dispatch<uint16_t>(d, syn, w, (uint16_t*)out, pos, offsets, overflow, cap, n, 0x3f80u, st);
    default:
      return write_dispatch<uint16_t>(d, syn, w, (uint16_t*)out, pos, offsets, overflow, cap, n, 0x3c00u, st);
  }
}

int tc_persp_write(const uint32_t* syn, float* out, uint8_t* pos, int64_t* offsets, int32_t* overflow,
                   int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n, void* stream) {
  return tc_persp_write_as(syn, out, TC_DTYPE_F32, pos, offsets, overflow, cap, workspace, workspace_bytes, d, n,
                           stream);
}

int tc_perspectives_as(const uint32_t* syn, void* out, int out_dtype, uint8_t* pos, int32_t* counts,
                       int64_t* offsets, int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes,
                       int d, int64_t n, void* stream) {
  if (out_dtype < TC_DTYPE_F32 || out_dtype > TC_DTYPE_F16) return -1;
  const int rc = tc_persp_count(syn, counts, offsets, workspace, workspace_bytes, d, n, stream);
  if (rc != 0) return rc;
  return tc_persp_write_as(syn, out, out_dtype, pos, offsets, overflow, cap, workspace, workspace_bytes, d, n,
                           stream);
}

int tc_perspectives(const uint32_t* syn, float* out, uint8_t* pos, int32_t* counts, int64_t* offsets,
                    int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n,
                    void* stream) {
  const int rc = tc_persp_count(syn, counts, offsets, workspace, workspace_bytes, d, n, stream);
  if (rc != 0) return rc;
  return tc_persp_write(syn, out, pos, offsets, overflow, cap, workspace, workspace_bytes, d, n, stream);
}

// tc_perspectives for callers that have no device allocator to take a workspace from (see the header).  The count,
// scan and writer kernels are the ones tc_perspectives launches; only the workspace differs: it is the current
// device's g_static_scratch.  Nothing in it has to be cleared between calls -- local_off, block_sum, block_off and
// chunk_sum are each rewritten in full for (d, n) by the count / scan kernels of this call before the writer reads
// them, and the look-back descriptors of the single-pass step (the one region that must start at zero) are not used
// on this path.  No cudaMalloc / cudaFree here or below: the set-up is cudaGetSymbolAddress, and the writer's
// once-per-device cudaFuncSetAttribute / occupancy query.
int tc_perspectives_static(const uint32_t* syn, float* out, uint8_t* pos, int32_t* counts, int64_t* offsets,
                           int32_t* overflow, int64_t cap, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  const size_t need = carve(nullptr, d, n).bytes;
  if (need > (size_t)TC_STATIC_SCRATCH_BYTES) {
    fprintf(stderr,
            "tc_perspectives_static: %lld envs of size %d need %zu bytes of scratch, the built-in array holds %zu "
            "(TC_STATIC_SCRATCH_BYTES); split the batch or use tc_perspectives with a caller-owned workspace\n",
            (long long)n, d, need, (size_t)TC_STATIC_SCRATCH_BYTES);
    return -2;
  }
  void* scratch = nullptr;
  const int rc = static_scratch(&scratch);
  if (rc != 0) return rc;
  return tc_perspectives(syn, out, pos, counts, offsets, overflow, cap, scratch, (size_t)TC_STATIC_SCRATCH_BYTES, d, n,
                         stream);
}

int tc_actor_pre(uint32_t* qubits, uint32_t* syn, uint32_t* prev_syn, int32_t* actions, float* reward,
                 uint8_t* terminal, uint8_t* gs, uint8_t* status, int32_t* steps, uint32_t* episode,
                 const uint32_t* p_thresh, int64_t* counters, int flags, int max_steps, int mode, int n_fixed,
                 uint64_t seed, int64_t env_base, int32_t* counts, int64_t* offsets, void* workspace,
                 size_t workspace_bytes, int d, int64_t n, void* stream) {
  return actor_pre_impl(qubits, syn, prev_syn, actions, reward, terminal, gs, status, steps, episode, p_thresh, counters,
                        flags, max_steps, mode, n_fixed, seed, env_base, counts, offsets, workspace, workspace_bytes, d,
                        n, stream, nullptr);
}

int tc_actor_step(uint32_t* qubits, uint32_t* syn, uint32_t* prev_syn, int32_t* actions, float* reward,
                  uint8_t* terminal, uint8_t* gs, uint8_t* status, int32_t* steps, uint32_t* episode,
                  const uint32_t* p_thresh, int64_t* counters, int flags, int max_steps, int mode, int n_fixed,
                  uint64_t seed, int64_t env_base, float* out, uint8_t* pos, int32_t* counts, int64_t* offsets,
                  int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n,
                  void* stream) {
  if (flags & TC_ACTOR_SINGLE_PASS) {
    if (bad_d(d) || d > TC_SINGLE_PASS_MAX_D || n < 0 || cap < 0 || !offsets) return -1;
    if (mode < 0 || mode > 2 || n_fixed < 0 || (mode != 0 && n_fixed == 0)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
      cudaMemsetAsync(offsets, 0, sizeof(int64_t), st);
      return launch_status();
    }
    if (!qubits || !syn || !actions || !reward || !terminal || !gs || !steps || !episode || !p_thresh || !counts ||
        !workspace || (cap > 0 && !out))
      return -1;
    if (((uintptr_t)actions & 15) || ((uintptr_t)out & 15) || ((uintptr_t)pos & 3)) return -1;
    const Workspace w = carve(workspace, d, n);
    if (workspace_bytes < w.bytes) return -2;
    ActorArgs a;
    a.qubits = qubits; a.syn = syn; a.prev_syn = prev_syn; a.actions = actions; a.reward = reward;
    a.terminal = terminal; a.gs = gs; a.status = status; a.steps = steps; a.episode = episode;
    a.p_thresh = p_thresh; a.counters = counters; a.counts = counts; a.local_off = w.local_off;
    a.block_sum = w.block_sum; a.key = make_key(seed); a.env_base = env_base; a.n = n; a.flags = flags;
    a.max_steps = max_steps; a.mode = mode; a.n_fixed = n_fixed; a.d = d; a.state_rows = nullptr;
    return fused_dispatch(a, w, out, pos, offsets, overflow, cap, st);
  }
  const int rc = tc_actor_pre(qubits, syn, prev_syn, actions, reward, terminal, gs, status, steps, episode,
                              p_thresh, counters, flags, max_steps, mode, n_fixed, seed, env_base, counts, offsets,
                              workspace, workspace_bytes, d, n, stream);
  if (rc != 0) return rc;
  return tc_persp_write(syn, out, pos, offsets, overflow, cap, workspace, workspace_bytes, d, n, stream);
}

int tc_host_actor_step(const tc_host_step* s) {
  if (!s || bad_d(s->d) || s->n < 0) return -1;
  if (s->state_mode < 0 || s->state_mode > 2) return -1;
  if (s->n == 0) return 0;
  // (a NULL stream handle is the legacy default stream, which is what torch's default current stream is)
  if (!s->h_actions || !s->d_actions || !s->d_stage || !s->h_stage || !s->ev_actions || !s->ev_compute ||
      !s->ev_result || (s->state_mode && !s->ev_unpacked))
    return -1;
  if ((s->o_reward & 3) || (s->o_counts & 3) || (s->o_state & 15)) return -1;
  cudaStream_t h2d = (cudaStream_t)s->h2d_stream, cs = (cudaStream_t)s->compute_stream, d2h = (cudaStream_t)s->d2h_stream;
  cudaMemcpyAsync(s->d_actions, s->h_actions, (size_t)s->n * 16, cudaMemcpyHostToDevice, h2d);
  cudaEventRecord((cudaEvent_t)s->ev_actions, h2d);
  cudaStreamWaitEvent(cs, (cudaEvent_t)s->ev_actions, 0);
  if (s->ev_unpacked_prev) cudaStreamWaitEvent(cs, (cudaEvent_t)s->ev_unpacked_prev, 0);
  float* reward = reinterpret_cast<float*>(s->d_stage + s->o_reward);
  int32_t* counts = reinterpret_cast<int32_t*>(s->d_stage + s->o_counts);
  uint8_t* terminal = s->d_stage + s->o_term;
  // state_mode 2: the step kernel itself writes the row bitmasks into the download block -- no second kernel has to
  // find SM slots underneath the perspective writer before the results can leave
  int rc = actor_pre_impl(s->qubits, s->syn, s->prev_syn, s->d_actions, reward, terminal, s->gs, s->status, s->steps,
                          s->episode, s->p_thresh, s->counters, s->flags, s->max_steps, s->mode, s->n_fixed, s->seed,
                          s->env_base, counts, s->offsets, s->workspace, s->workspace_bytes, s->d, s->n, cs,
                          s->state_mode == 2 ? s->d_stage + s->o_state : nullptr);
  if (rc != 0) return rc;
  cudaEventRecord((cudaEvent_t)s->ev_compute, cs);
  rc = tc_persp_write_as(s->syn, s->out, s->out_dtype, s->pos, s->offsets, s->overflow, s->cap, s->workspace,
                         s->workspace_bytes, s->d, s->n, cs);
  if (rc != 0) return rc;
  cudaStreamWaitEvent(d2h, (cudaEvent_t)s->ev_compute, 0);
  if (s->state_mode == 1) {
    rc = tc_unpack_state(s->syn, s->d_stage + s->o_state, s->d, s->n, d2h);
    if (rc != 0) return rc;
  }
  if (s->state_mode) cudaEventRecord((cudaEvent_t)s->ev_unpacked, d2h);
  cudaMemcpyAsync(s->h_stage, s->d_stage, s->stage_bytes, cudaMemcpyDeviceToHost, d2h);
  cudaEventRecord((cudaEvent_t)s->ev_result, d2h);
  return launch_status();
}

// ---- policy-side entry points ---------------------------------------------

int tc_select_action(const float* q_values, const uint8_t* pos, const int64_t* offsets, const uint32_t* episode,
                     const int32_t* steps, const uint32_t* eps_thresh, uint32_t eps_scalar,
                     const uint8_t* greedy_in, uint64_t seed, int64_t env_base, int32_t* actions, float* q_sel,
                     int64_t* sel_idx, uint8_t* was_greedy, int64_t n, void* stream) {
  if (n < 0) return -1;
  if (n == 0) return 0;
  if (!q_values || !pos || !offsets || !actions || !q_sel) return -1;
  if (((uintptr_t)actions & 15) || ((uintptr_t)pos & 3)) return -1;
  select_action_kernel<<<grid_for(n * 32, 256), 256, 0, (cudaStream_t)stream>>>(
      q_values, pos, offsets, episode, steps, eps_thresh, eps_scalar, greedy_in, make_key(seed), env_base, actions,
      q_sel, sel_idx, was_greedy, n);
  return launch_status();
}

int tc_p_error_curriculum(const int32_t* steps, const uint32_t* episode, uint32_t* p_cap, uint32_t* p_thresh,
                           uint32_t start_thresh, uint32_t delta_thresh, uint32_t final_thresh, int random_strategy,
                           int all_envs, uint64_t seed, int64_t env_base, int64_t n, void* stream) {
  if (n < 0) return -1;
  if (n == 0) return 0;
  if (!steps || !episode || !p_cap || !p_thresh) return -1;
  p_error_curriculum_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(
      steps, episode, p_cap, p_thresh, start_thresh, delta_thresh, final_thresh, random_strategy, all_envs,
      make_key(seed), env_base, n);
  return launch_status();
}

int tc_transitions(const uint32_t* syn_prev, const uint32_t* syn_next, const int32_t* actions, void* out_prev,
                   void* out_next, int32_t* action_out, int out_dtype, int d, int64_t n, void* stream) {
  if (bad_d(d) || n < 0) return -1;
  if (out_dtype != TC_DTYPE_F32 && out_dtype != TC_DTYPE_U8) return -1;
  if (n == 0) return 0;
  if (!actions) return -1;
  if ((out_prev && !syn_prev) || (out_next && !syn_next)) return -1;
  if (((uintptr_t)actions & 15) || ((uintptr_t)action_out & 15)) return -1;
  const unsigned grid = grid_for(n * 2 * d * d, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == TC_DTYPE_F32)
    transition_kernel<float><<<grid, 256, 0, st>>>(syn_prev, syn_next, actions, (float*)out_prev, (float*)out_next,
                                                   action_out, d, n);
  else
    transition_kernel<uint8_t><<<grid, 256, 0, st>>>(syn_prev, syn_next, actions, (uint8_t*)out_prev,
                                                     (uint8_t*)out_next, action_out, d, n);
  return launch_status();
}

int tc_td_priorities(const int32_t* actions, const float* reward, const float* q, const float* q_next,
                     float discount, float* priorities, int64_t n, void* stream) {
  if (n < 0) return -1;
  if (n == 0) return 0;
  if (!actions || !reward || !q || !q_next || !priorities) return -1;
  td_priority_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(actions, reward, q, q_next, discount,
                                                                        priorities, n);
  return launch_status();
}

int tc_segmented_max(const float* q_values, const int64_t* offsets, int pad_to, const int32_t* pad_to_dev,
                     float* out, int64_t n, void* stream) {
  if (n < 0) return -1;
  if (n == 0) return 0;
  if (!q_values || !offsets || !out) return -1;
  segmented_max_kernel<<<grid_for(n * 32, 256), 256, 0, (cudaStream_t)stream>>>(q_values, offsets, pad_to,
                                                                               pad_to_dev, out, n);
  return launch_status();
}

}  // extern "C"
