// Single-pass actor step: the step / bookkeeping / auto-reset of actor_step_kernel, the cross-env scan of
// scan_blocks_kernel and the perspective writer of persp_write_kernel in ONE kernel (SURVEY.md section 7 step 5:
// "the fused single-pass variant (decoupled look-back)").
//
// One env per warp, CTAs in env order.  A warp
//   1. steps its env with the same device routines the three-kernel path uses (lane r holds row r: group size 32),
//      stores the new state and results, and counts the env's perspectives P;
//   2. PUBLISHES P:  desc0[e] = epoch<<12 | P                                   (one relaxed 32-bit store)
//   3. obtains its exclusive offset by decoupled look-back (Merrill & Garland 2016) on two levels:
//        inside its group of 32 consecutive envs it reads the 32 desc0 words with one coalesced load and sums the
//        predecessors' P; the warp of the group's last env ("closer") publishes the group aggregate
//        desc1[g] = {epoch, AGGREGATE, sum} as soon as it has it, looks back over the earlier groups' descriptors
//        32 at a time (nearest {PREFIX} + the aggregates after it) and then publishes {epoch, PREFIX, inclusive
//        sum}.  Every other warp does the same look-back for itself, so nobody waits for a closer's look-back --
//        only for aggregates, which are published without waiting for anything;
//   4. streams its tiles exactly like persp_write_kernel (write_env).
// Forward progress: descriptors only ever wait for envs with a LOWER index, CTAs are dispatched in index order and a
// publish never waits, so the lowest unpublished env can always make progress (the argument cub::DeviceScan relies
// on).  Descriptors carry a 20-bit epoch instead of being cleared: the header's launch counter is bumped by the
// warp of the last env once its look-back is done (every other env has published, hence read the counter, by
// then); a programmatically dependent successor reads it only after this grid has completed (pdl_wait).  The
// workspace must be zero-filled once before its first use.
#pragma once
#include "persp_write.cuh"
#include "state_kernels.cuh"

namespace tc {

struct ActorArgs {
  uint32_t* qubits;
  uint32_t* syn;
  uint32_t* prev_syn;
  int32_t* actions;
  float* reward;
  uint8_t* terminal;
  uint8_t* gs;
  uint8_t* status;
  int32_t* steps;
  uint32_t* episode;
  const uint32_t* p_thresh;
  int64_t* counters;
  int32_t* counts;
  int32_t* local_off;
  int32_t* block_sum;
  uint2 key;
  int64_t env_base;
  int64_t n;
  int flags, max_steps, mode, n_fixed, d;
  uint8_t* state_rows;   // nullable: the new states as lattice-row bitmasks (tc_state_rows layout), written by the step
};

constexpr uint32_t kDescAggregate = 1, kDescPrefix = 2;

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// group descriptor: epoch (20 bits) | state (2 bits) | value (42 bits)
__device__ __forceinline__ unsigned long long pack_desc1(uint32_t epoch, uint32_t state, long long value) {
  return ((unsigned long long)epoch << 44) | ((unsigned long long)state << 42) | (unsigned long long)value;
}

__device__ __forceinline__ long long wsum64(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct ScanWs {
  uint32_t* header;                // [16]: word 0 = launch counter
  uint32_t* desc0;                 // [n]
  unsigned long long* desc1;       // [ceil(n/32)]
};

// Exclusive offset of env e (perspective count P already published in desc0[e]); all lanes return the same value.
__device__ __forceinline__ long long lookback_offset(const ScanWs& ws, uint32_t epoch, int64_t e, int P, int lane) {
  constexpr uint32_t kFull = 0xffffffffu;
  const int64_t g = e >> 5;
  const int idx = (int)(e & 31);
  // ---- predecessors inside the group ----
  uint32_t dv = epoch << 12;
  if (lane < idx) {
    const uint32_t* p = ws.desc0 + (g << 5) + lane;
    dv = ld_relaxed_u32(p);
    while ((dv >> 12) != epoch) {
      __nanosleep(40);
      dv = ld_relaxed_u32(p);
    }
  }
  __syncwarp();
  const int sum_in = gsum<32>(kFull, (int)(dv & 0xfffu));       // lanes >= idx contribute 0
  const bool closer = idx == 31;
  if (closer && lane == 0) st_relaxed_u64(ws.desc1 + g, pack_desc1(epoch, kDescAggregate, sum_in + P));
  // ---- earlier groups: nearest inclusive prefix + the aggregates after it ----
  long long base = 0;
  int64_t hi = g - 1;                      // nearest group not yet accounted for
  while (hi >= 0) {
    const int64_t gi = hi - lane;
    unsigned long long d1 = pack_desc1(epoch, kDescPrefix, 0);   // virtual group -1: inclusive prefix 0
    bool done = false;
    while (true) {
      if (gi >= 0) d1 = ld_relaxed_u64(ws.desc1 + gi);
      const bool ready = (uint32_t)(d1 >> 44) == epoch;
      const uint32_t rmask = __ballot_sync(kFull, ready);
      const uint32_t pmask = __ballot_sync(kFull, ready && ((d1 >> 42) & 3ull) == kDescPrefix);
      if (pmask) {
        const int first = __ffs(pmask) - 1;                      // nearest group holding an inclusive prefix
        const uint32_t below = first == 0 ? 0u : (0xffffffffu >> (32 - first));
        if ((rmask & below) == below) {
          base += wsum64(lane <= first ? (long long)(d1 & ((1ull << 42) - 1ull)) : 0ll);
          done = true;
          break;
        }
      } else if (rmask == kFull) {
        base += wsum64((long long)(d1 & ((1ull << 42) - 1ull)));
        break;
      }
      __nanosleep(60);
    }
    if (done) break;
    hi -= 32;
  }
  if (closer && lane == 0) st_relaxed_u64(ws.desc1 + g, pack_desc1(epoch, kDescPrefix, base + sum_in + P));
  return base + sum_in;
}

template <int D, typename OutT>
__global__ void __launch_bounds__(PerspCfg<D>::kWarps * 32)
fused_step_kernel(const ActorArgs a, const ScanWs ws, OutT* __restrict__ out, uint8_t* __restrict__ pos,
                  int64_t* __restrict__ offsets, int32_t* __restrict__ overflow, int64_t cap, uint32_t one16) {
  using Cfg = PerspCfg<D>;
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr int G = 32;                           // one env per warp: lane r holds row r
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const WriterCtx<D, OutT> cx(smem, warp, lane);
  build_cand_table<D>(smem, Cfg::kWarps * 32);
  pdl_wait();                 // the previous step (same workspace, same state buffers) must have completed
  pdl_launch_dependents();
  const int64_t e = (int64_t)blockIdx.x * Cfg::kWarps + warp;
  uint32_t ring = 0;
  if (e < a.n) {                                  // warp-uniform
#ifdef TC_FUSED_TIMING
    unsigned long long t_start;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
    const uint32_t launch = *ws.header;
    const uint32_t epoch = launch % 0xfffffu + 1u;             // never 0: zero-filled descriptors are "not ready"
    const int r = lane;
    const bool lead = lane == 0;
    Rows s;
    load_qubits<G>(a.qubits, e, D, r, s);
    load_syn<G>(a.syn, e, D, r, s);
    uint32_t ep = a.episode[e];
    int stp = a.steps[e];
    const uint64_t env_id = (uint64_t)(a.env_base + e);
    int am, ai, aj, aop;
    if (a.flags & TC_ACTOR_RANDOM_ACTIONS) {
      uint32_t q0, q1;
      persp_masks<G>(kFull, D, r, s.v, s.pl, q0, q1);
      random_action<G>(kFull, D, r, q0, q1, env_id, ep, (uint32_t)stp, a.key, am, ai, aj, aop);
      if (lead) *reinterpret_cast<int4*>(a.actions + e * 4) = make_int4(am, ai, aj, aop);
    } else {
      const int4 act = *reinterpret_cast<const int4*>(a.actions + e * 4);
      am = act.x; ai = act.y; aj = act.z; aop = act.w;
    }
    const int before = gsum<G>(kFull, __popc(s.v) + __popc(s.pl));
    const bool frozen = (a.flags & TC_ACTOR_FREEZE_DONE) && before == 0;
    const bool ok = frozen ? true : apply_action<G>(D, r, am, ai, aj, aop, s);
    const int after = gsum<G>(kFull, __popc(s.v) + __popc(s.pl));
    const bool term = after == 0;
    const bool even = parity_even<G>(kFull, s);
    if (!frozen) stp += 1;
    unsigned long long* ctr = reinterpret_cast<unsigned long long*>(a.counters);
    if (lead && !frozen) {
      a.reward[e] = term ? 100.0f : (float)(before - after);
      a.terminal[e] = term ? 1 : 0;
      a.gs[e] = even ? 1 : 0;
      if (ctr && (a.flags & TC_ACTOR_FREEZE_DONE)) atomicAdd(ctr + kCtrEnvSteps, 1ull);   // else: n, added once below
      if (!ok) {
        if (a.status) a.status[e] |= kStatusBadAction;
        if (ctr) atomicAdd(ctr + kCtrBadActions, 1ull);
      }
    }
    if (a.prev_syn) store_syn<G>(a.prev_syn, e, D, r, s);
    const bool done = !frozen && (term || stp > a.max_steps);      // Actor_mp.py:171-172
    if ((a.flags & TC_ACTOR_AUTO_RESET) && done) {
      uint8_t st = 0;
      SampleStats stats;
      sample_errors<G>(kFull, D, r, env_id, ep, a.p_thresh[e], a.mode, a.n_fixed, a.key, s, st, stats);
      const int nx = gsum<G>(kFull, stats.nx), ny = gsum<G>(kFull, stats.ny), nz = gsum<G>(kFull, stats.nz);
      ep += 1;
      stp = 0;
      if (lead) {
        if (ctr) {
          atomicAdd(ctr + kCtrEpisodes, 1ull);
          atomicAdd(ctr + (term ? kCtrSolved : kCtrTimeouts), 1ull);
          if (term && even) atomicAdd(ctr + kCtrGroundKept, 1ull);
          atomicAdd(ctr + kCtrErrX, (unsigned long long)nx);
          atomicAdd(ctr + kCtrErrY, (unsigned long long)ny);
          atomicAdd(ctr + kCtrErrZ, (unsigned long long)nz);
          atomicAdd(ctr + kCtrQubitsSampled, (unsigned long long)(2 * D * D));
          atomicAdd(ctr + kCtrResetAttempts, (unsigned long long)stats.attempts);
        }
        if (st && a.status) a.status[e] |= st;
      }
    }
    store_qubits<G>(a.qubits, e, D, r, s);
    store_syn<G>(a.syn, e, D, r, s);
    uint32_t m0, m1;
    persp_masks<G>(kFull, D, r, s.v, s.pl, m0, m1);
    const int P = gsum<G>(kFull, __popc(m0) + __popc(m1));
    if (lead) {
      st_relaxed_u32(ws.desc0 + e, (epoch << 12) | (uint32_t)P);   // publish first: others wait for this
      a.steps[e] = stp;
      a.episode[e] = ep;
      a.counts[e] = P;
    }
#ifdef TC_FUSED_TIMING   /* diagnostic build: per-env phase durations (ns >> 4, 16 bits each) into the unused local_off */
    unsigned long long t_pub;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_pub));
#endif
#ifdef TC_FUSED_NO_LOOKBACK   /* diagnostic build: no scan at all, every env writes at its worst-case slot */
    const long long off = (long long)e * (2 * D * D);
#else
    const long long off = lookback_offset(ws, epoch, e, P, lane);
#endif
#ifdef TC_FUSED_TIMING
    {
      unsigned long long t_done;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_done));
      const unsigned long long pre = (t_pub - t_start) >> 4, wait = (t_done - t_pub) >> 4;
      if (lead) a.local_off[e] = (int)((pre > 0xffffull ? 0xffffull : pre) | ((wait > 0xffffull ? 0xffffull : wait) << 16));
    }
#endif
    if (lead) offsets[e] = off;
    if (e == a.n - 1 && lead) {                    // every other env has published (and read the launch counter)
      offsets[a.n] = off + P;
      *ws.header = launch + 1u;
      if (ctr) {
        atomicAdd(ctr + kCtrPerspectives, (unsigned long long)(off + P));
        if (!(a.flags & TC_ACTOR_FREEZE_DONE)) atomicAdd(ctr + kCtrEnvSteps, (unsigned long long)a.n);
      }
    }
    if (P > 0) {
      if (off + P > cap) {
        if (lead && overflow) *overflow = 1;
      } else {
        write_env<D, OutT>(cx, s.v, s.pl, m0, m1, P, off, 0, 1, ring, out, pos, one16);
      }
    }
  }
  __syncwarp();
  if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

}  // namespace tc
