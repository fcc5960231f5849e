// Policy-side kernels around the env hot path:
//   select_action_kernel   _selectActionBatch_prime + epsilon-greedy draw (src/numba/util_actor.py:48-107,
//                          single-env form src/util_actor.py:47-65)
//   transition_kernel      generateTransitionParallel's state pair (src/util_actor.py:223-264):
//                          shift_state (src/util.py:97-102) + rotate_state (src/util.py:87-94)
//   td_priority_kernel     computePrioritiesParallel (src/util_actor.py:268-287)
//   segmented_max_kernel   the per-state max of predictMaxOptimized (src/util_learner.py:88-111),
//                          including its zero-padding quirk
#pragma once
#include <cfloat>

#include "common.cuh"

namespace tc {

// kDomSelect (common.cuh): Philox counter word 3 domain of the epsilon-greedy draws

// One warp per env.  Q-values of the env are the rows offsets[e] .. offsets[e+1]-1 of q (float32 [sumP][3]).
// Greedy: first maximum in row-major (perspective, action) order == np.argmax / np.where(...)[0]
// (src/util_actor.py:51, src/numba/util_actor.py:93-95).  Random: perspective mulhi(u1, P), action mulhi(u2, 3).
// The draw is Philox(env id, episode, kDomSelect | step): random iff u0 < eps_thresh (P(random) = epsilon,
// the reference's `(1 - epsilon) > rand`), unless greedy_in overrides the decision.
// Envs without perspectives (terminal state) get the identity action (0,0,0,0): rule_table[c,0] = c.
__global__ void __launch_bounds__(256)
select_action_kernel(const float* __restrict__ q, const uint8_t* __restrict__ pos,
                     const int64_t* __restrict__ offsets, const uint32_t* __restrict__ episode,
                     const int32_t* __restrict__ steps, const uint32_t* __restrict__ eps_thresh,
                     uint32_t eps_scalar, const uint8_t* __restrict__ greedy_in, uint2 key, int64_t env_base,
                     int32_t* __restrict__ actions, float* __restrict__ q_sel, int64_t* __restrict__ sel_idx,
                     uint8_t* __restrict__ was_greedy, int64_t n) {
  const int lane = threadIdx.x & 31;
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= n) return;
  const int64_t off = offsets[e];
  const int P = (int)(offsets[e + 1] - off);
  if (P <= 0) {
    if (lane == 0) {
      *reinterpret_cast<int4*>(actions + e * 4) = make_int4(0, 0, 0, 0);
      q_sel[e * 3] = q_sel[e * 3 + 1] = q_sel[e * 3 + 2] = 0.0f;
      if (sel_idx) sel_idx[e] = -1;
      if (was_greedy) was_greedy[e] = 1;
    }
    return;
  }
  const uint64_t env_id = (uint64_t)(env_base + e);
  const uint32_t ep = episode ? episode[e] : 0u;
  const uint32_t stp = steps ? (uint32_t)steps[e] : 0u;
  const uint4 rnd = philox4x32_10(make_uint4((uint32_t)env_id, (uint32_t)(env_id >> 32), ep, kDomSelect | (stp & kSelectStepMask)), key);
  const uint32_t thr = eps_thresh ? eps_thresh[e] : eps_scalar;
  const bool greedy = greedy_in ? (greedy_in[e] != 0) : !(rnd.x < thr);
  int p, a;
  if (greedy) {
    const float* qe = q + off * 3;
    float best = -FLT_MAX;
    int bi = 0x7fffffff;
    bool seen = false;
    for (int t = lane; t < 3 * P; t += 32) {
      const float v = __ldg(qe + t);
      if (!seen || v > best) {      // strict: the earliest index wins inside a lane
        best = v;
        bi = t;
        seen = true;
      }
    }
    if (!seen) best = -FLT_MAX;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > best || (ov == best && oi < bi))) {
        best = ov;
        bi = oi;
      }
    }
    if (bi == 0x7fffffff) bi = 0;
    p = bi / 3;
    a = bi - p * 3;
  } else {
    p = (int)mulhi32(rnd.y, (uint32_t)P);
    a = (int)mulhi32(rnd.z, 3u);
  }
  if (lane == 0) {
    const uint32_t pw = reinterpret_cast<const uint32_t*>(pos)[off + p];
    *reinterpret_cast<int4*>(actions + e * 4) =
        make_int4((int)(pw & 0xffu), (int)((pw >> 8) & 0xffu), (int)((pw >> 16) & 0xffu), a + 1);
    if (sel_idx) sel_idx[e] = off + p;
    if (was_greedy) was_greedy[e] = greedy ? 1 : 0;
  }
  if (lane < 3) q_sel[e * 3 + lane] = __ldg(q + (off + p) * 3 + lane);
}

// One thread per output element.  View of state S centred on the acted qubit (m,i,j), g = d/2:
//   m=0: out[c,a,b] = S[c][(a+i-g)%d][(b+j-g)%d]                                  (shift_state)
//   m=1: out[0,a,b] = V [(b+i-g)%d][((d-a)%d + j-g)%d]                             (shift then rotate_state)
//        out[1,a,b] = Pl[(b+i-g)%d][(d-1-a + j-g)%d]
// Written for both the pre-step and the post-step syndrome; the stored action becomes (m,g,g,op).
template <typename OutT>
__global__ void __launch_bounds__(256)
transition_kernel(const uint32_t* __restrict__ syn_prev, const uint32_t* __restrict__ syn_next,
                  const int32_t* __restrict__ actions, OutT* __restrict__ out_prev, OutT* __restrict__ out_next,
                  int32_t* __restrict__ action_out, int d, int64_t n) {
  const int per = 2 * d * d;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * per) return;
  const int64_t e = t / per;
  const int idx = (int)(t - e * per);
  const int c = idx / (d * d);
  const int rem = idx - c * d * d;
  const int a = rem / d, b = rem - a * d;
  const int4 act = *reinterpret_cast<const int4*>(actions + e * 4);
  const int g = d / 2;
  const int m = act.x & 1;
  const int i = ((act.y % d) + d) % d, j = ((act.z % d) + d) % d;
  int row, col;
  if (m == 0) {
    row = (a + i - g + d) % d;
    col = (b + j - g + d) % d;
  } else {
    row = (b + i - g + d) % d;
    col = c == 0 ? ((d - a) % d + j - g + d) % d : (d - 1 - a + j - g + d) % d;
  }
  const int64_t w = e * 2 * d + c * d + row;
  if (out_prev) out_prev[t] = (OutT)((syn_prev[w] >> col) & 1u);
  if (out_next) out_next[t] = (OutT)((syn_next[w] >> col) & 1u);
  if (action_out && idx == 0) *reinterpret_cast<int4*>(action_out + e * 4) = make_int4(m, g, g, act.w);
}

// computePrioritiesParallel (src/util_actor.py:268-287): |R + discount * max_a Qns[a] - Q[action-1]|.
__global__ void __launch_bounds__(256)
td_priority_kernel(const int32_t* __restrict__ actions, const float* __restrict__ reward,
                   const float* __restrict__ q, const float* __restrict__ q_next, float discount,
                   float* __restrict__ prio, int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const int a = actions[e * 4 + 3] - 1;
  const float qn = fmaxf(fmaxf(q_next[e * 3], q_next[e * 3 + 1]), q_next[e * 3 + 2]);
  const float qv = q[e * 3 + (a < 0 ? 0 : (a > 2 ? 2 : a))];
  prio[e] = fabsf(reward[e] + discount * qn - qv);
}

// predictMaxOptimized's reduction (src/util_learner.py:96-111): per state the max over its
// (perspective, action) Q-values; states with fewer perspectives than the batch maximum are zero-padded
// in the reference (:100-106), so 0 takes part in their max; terminal states (no perspectives) give 0.
__global__ void __launch_bounds__(256)
segmented_max_kernel(const float* __restrict__ q, const int64_t* __restrict__ offsets, int pad_to,
                     const int32_t* __restrict__ pad_to_dev, float* __restrict__ out, int64_t n) {
  const int lane = threadIdx.x & 31;
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (e >= n) return;
  const int64_t off = offsets[e];
  const int P = (int)(offsets[e + 1] - off);
  float best = -FLT_MAX;
  for (int t = lane; t < 3 * P; t += 32) best = fmaxf(best, __ldg(q + off * 3 + t));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
  const int pad = pad_to_dev ? *pad_to_dev : pad_to;   // largest_persp_batch (src/util_learner.py:83)
  if (P < pad || P == 0) best = fmaxf(best, 0.0f);
  if (lane == 0) out[e] = best;
}

// The p_error curriculum of the reference's actor (src/Actor_mp.py:171-181; parameters Distributed_mp.py:53-56,186-189)
// in threshold units T = floor(p * 2^32): on every reset the env's ceiling grows, cap <- min(final, cap + delta), and
// the p of the new episode is the ceiling ('linear') or uniform on [start, cap] ('random').  One thread per env.  Run
// right after the fused step: an env whose steps[e] == 0 was auto-reset in that step and now receives the ceiling and
// the p_thresh of its NEXT reset (all != 0: every env, to prime the first one) -- the k-th reset of an env therefore
// uses exactly the reference's k-th ceiling, drawn one episode ahead of time.  The 'random' draw is
// Philox(env id, episode[e], kDomPError).x scaled to the integer interval [min(start, cap), cap].
__global__ void __launch_bounds__(256)
p_error_curriculum_kernel(const int32_t* __restrict__ steps, const uint32_t* __restrict__ episode,
                          uint32_t* __restrict__ p_cap, uint32_t* __restrict__ p_thresh, uint32_t start_t,
                          uint32_t delta_t, uint32_t final_t, int random, int all, uint2 key, int64_t env_base,
                          int64_t n) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  if (!all && steps[e] != 0) return;
  const unsigned long long grown = (unsigned long long)p_cap[e] + (unsigned long long)delta_t;
  const uint32_t cap = grown < (unsigned long long)final_t ? (uint32_t)grown : final_t;
  p_cap[e] = cap;
  uint32_t p = cap;
  if (random) {
    const uint32_t lo = start_t < cap ? start_t : cap;
    const uint64_t env_id = (uint64_t)(env_base + e);
    const uint4 rnd = philox4x32_10(make_uint4((uint32_t)env_id, (uint32_t)(env_id >> 32), episode[e], kDomPError), key);
    p = lo + (uint32_t)(((unsigned long long)rnd.x * ((unsigned long long)(cap - lo) + 1ull)) >> 32);
  }
  p_thresh[e] = p;
}

}  // namespace tc
