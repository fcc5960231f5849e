// Shared device helpers for the toric-code environment engine (sm_100a).
//
// HBM layout (see DESIGN.md "Data layout"):
//   qubits : uint32 [N][4][d]   planes x0, x1, z0, z1; one lattice row per 32-bit word, bit j = column j
//   syn    : uint32 [N][2][d]   planes V (vertex), Pl (plaquette); same row packing
// Pauli code c in {0:I,1:X,2:Y,3:Z} (reference gym_ToricCode/.../ToricCode.py:57) is stored as the two
// bit planes x = [c in {1,2}], z = [c in {2,3}], which turns the reference's rule_table (ToricCode.py:57,
// :79-81) into XOR on the planes.
//
// State kernels use one sub-warp "group" of G lanes per environment with lane r holding row r of every
// plane (G = 8, 16 or 32 for d <= 8, 16, 32); row neighbours are group shuffles, column neighbours are
// rotations inside the d-bit row word.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace tc {

constexpr int kMinD = 3;
constexpr int kMaxD = 32;

// status bits (per env, sticky until cleared by the host mirror)
constexpr uint8_t kStatusBadAction = 1;     // action outside the lattice / op > 3: masked to a no-op
constexpr uint8_t kStatusResetGaveUp = 2;   // rejection sampler hit its attempt cap (p_error ~ 0)

// counters[] slots (int64) accumulated by the fused actor step
enum Counter : int {
  kCtrEnvSteps = 0,
  kCtrEpisodes = 1,
  kCtrSolved = 2,
  kCtrGroundKept = 3,   // solved AND no odd plane parity at episode end (evalGroundState, ToricCode.py:326-335)
  kCtrTimeouts = 4,
  kCtrPerspectives = 5,
  kCtrErrX = 6,
  kCtrErrY = 7,
  kCtrErrZ = 8,
  kCtrQubitsSampled = 9,
  kCtrResetAttempts = 10,
  kCtrBadActions = 11,
  kNumCounters = 16
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11), counter-based: the draw for (env, episode, slot) does not
// depend on which GPU or thread computes it.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t mulhi32(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
  return __umulhi(a, b);
#else
  return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

__host__ __device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = mulhi32(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = mulhi32(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

// Counter word 3 domains (word 0/1 = global env id, word 2 = episode number) -- ALL of them, so that two kinds of
// draw of the same (env, episode) can never land on the same Philox block.  A domain is a tag in the top bits plus a
// payload that stays below its tag's lowest bit:
//   field   0x00000000 | attempt<<10 | row<<5 | call     attempt < 2^20 -> payload < 2^30          (sampler.cuh)
//   first   0x40000000 | attempt                         attempt < 2^20 -> payload < 2^29          (sampler.cuh)
//   select  0x60000000 | step                            step masked to 29 bits                    (policy_kernels.cuh)
//   fixed   0x80000000 | attempt<<20 | draw              attempt < 2^9, draw < 2^20 -> < 2^30      (sampler.cuh)
//   p-error 0xA0000000                                   no payload: one draw per (env, episode)   (policy_kernels.cuh)
//   action  0xC0000000 | step                            both tag bits set: OR-ing cannot leave it (state_kernels.cuh)
// (ABI 5 had select on 0x40000000 as well: the epsilon draw of (episode E, step 0) was the block of the first-error
// draw, attempt 0, of the conditioned reset that opens episode E -- an exploratory first action forced U < eps and
// with it a first error at a low linear index.)
constexpr uint32_t kDomField = 0x00000000u;    // depolarizing field slots / code chunks
constexpr uint32_t kDomFirst = 0x40000000u;    // first-error draw of the conditioned sampler
constexpr uint32_t kDomSelect = 0x60000000u;   // epsilon-greedy draws of tc_select_action
constexpr uint32_t kDomFixed = 0x80000000u;    // fixed-weight sampler draws
constexpr uint32_t kDomPError = 0xA0000000u;   // p_error curriculum draw ('random' strategy) of tc_p_error_curriculum
constexpr uint32_t kDomAction = 0xC0000000u;   // random-policy action draws
constexpr uint32_t kSelectStepMask = 0x1fffffffu;
constexpr uint32_t kMaxAttempts = 1u << 20;    // depolarizing rejection attempts
// the fixed-weight counter carries 9 bits of attempt number ((attempt & 511) << 20): attempts past 512 would replay
// attempts 0..511 draw for draw and could never succeed
constexpr uint32_t kMaxFixedAttempts = 1u << 9;
static_assert(((kMaxAttempts - 1) << 10 | 0x3ffu) < kDomFirst, "field payload runs into the next domain");
static_assert((kDomFirst | (kMaxAttempts - 1)) < kDomSelect, "first-error payload runs into the select domain");
static_assert((kDomSelect | kSelectStepMask) < kDomFixed, "select payload runs into the fixed-weight domain");
static_assert((kDomFixed | ((kMaxFixedAttempts - 1) << 20) | 0xfffffu) < kDomPError, "fixed-weight payload too wide");
static_assert(kDomPError < kDomAction, "p-error domain below the action domain");
static_assert((kMaxFixedAttempts - 1) <= 511u, "the fixed-weight counter has 9 attempt bits");
static_assert((kDomAction >> 30) == 3u, "the action domain must own both tag bits (its step is OR-ed in unmasked)");
constexpr int kSampleOneDraw = 16;             // TC_SAMPLE_ONE_DRAW (OR-ed into the sampling mode)
// The depolarizing reset draws its field conditioned on >= 1 error while p * #qubits < kCondMeanErrors (sampler.cuh).
// At 8 an unconditioned draw is empty with probability < e^-8 = 3e-4, so a rejection pass -- which in the
// one-thread-per-env kernels stalls the whole warp -- is rare on both sides of the switch.
constexpr unsigned long long kCondMeanErrors = 8;

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch: the three kernels of one actor step (and the first kernel of the next
// step) are launched with cudaLaunchAttributeProgrammaticStreamSerialization, so a kernel's CTAs become
// resident while the previous kernel drains and only its reads of the predecessor's results wait.
// pdl_wait() returns once every prerequisite grid has completed and its writes are visible (a no-op for a
// normal launch); pdl_launch_dependents() lets the next kernel in the stream start launching.
//
// INVARIANT (every kernel launched through launch_pdl): NO EXIT PATH BEFORE pdl_wait().  Ordering across more than
// one launch is transitive only -- actor(k+1) waits for writer(k), which waited for scan(k), which waited for
// actor(k) -- and a grid only counts as complete for its dependents once all of its threads have exited; a thread
// that returned before its own pdl_wait() would let its grid complete while the grid in front of it is still
// running, and the dependent behind it would then race with that older grid on syn / workspace / offsets.  So an
// out-of-range thread clamps its index (or skips its work) and still executes pdl_wait().  Kernels launched with the
// plain <<<>>> syntax (reset_kernel, the policy kernels) take a full stream dependency and may return early.
// tests/test_gpu_actor.py::test_pdl_chain_matches_serialized_launches compares long chains against a TC_PDL=0 build.
// ---------------------------------------------------------------------------------------------
#ifndef TC_PDL
#define TC_PDL 1
#endif
__device__ __forceinline__ void pdl_wait() {
#if TC_PDL
  asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
__device__ __forceinline__ void pdl_launch_dependents() {
#if TC_PDL
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

// ---------------------------------------------------------------------------------------------
// d-bit row helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t row_mask(int d) { return d >= 32 ? 0xffffffffu : ((1u << d) - 1u); }
// bit j of result = bit (j-1) mod d of w   (value at column j-1 moves to column j)
__device__ __forceinline__ uint32_t rotl_d(uint32_t w, int d) {
  return ((w << 1) | (w >> (d - 1))) & row_mask(d);
}
// bit j of result = bit (j+1) mod d of w
__device__ __forceinline__ uint32_t rotr_d(uint32_t w, int d) {
  return ((w >> 1) | (w << (d - 1))) & row_mask(d);
}

// periodic neighbours without an integer division (d is a runtime value in the state kernels)
__device__ __forceinline__ int wrap_inc(int x, int d) { return x + 1 >= d ? x + 1 - d : x + 1; }
__device__ __forceinline__ int wrap_dec(int x, int d) { return x == 0 ? d - 1 : x - 1; }

template <int G>
__device__ __forceinline__ uint32_t group_mask() {
  if (G == 32) return 0xffffffffu;
  const unsigned lane = threadIdx.x & 31u;
  return ((1u << G) - 1u) << (lane & ~(unsigned)(G - 1));
}

template <int G>
__device__ __forceinline__ uint32_t gshfl(uint32_t gmask, uint32_t v, int src) {
  return __shfl_sync(gmask, v, src, G);
}

template <int G>
__device__ __forceinline__ int gsum(uint32_t gmask, int v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gmask, v, o, G);
  return v;
}

template <int G>
__device__ __forceinline__ uint32_t gxor(uint32_t gmask, uint32_t v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v ^= __shfl_xor_sync(gmask, v, o, G);
  return v;
}

// exclusive prefix sum across the group; total returned through `total`
template <int G>
__device__ __forceinline__ int gscan_excl(uint32_t gmask, int v, int lane_in_group, int& total) {
  int inc = v;
#pragma unroll
  for (int o = 1; o < G; o <<= 1) {
    const int t = __shfl_up_sync(gmask, inc, o, G);
    if (lane_in_group >= o) inc += t;
  }
  total = __shfl_sync(gmask, inc, G - 1, G);
  return inc - v;
}

// Rows of one environment held by one lane (row r = lane_in_group; lanes >= d hold zeros).
struct Rows {
  uint32_t x0, x1, z0, z1;   // qubit planes
  uint32_t v, pl;            // syndrome planes
};

// createSyndromOpt (ToricCode.py:168-199) as a bit stencil:
//   V[i,j]  = z0[i,j] ^ z0[i-1,j] ^ z1[i,j] ^ z1[i,j-1]
//   Pl[i,j] = x0[i,j] ^ x0[i,j+1] ^ x1[i,j] ^ x1[i+1,j]
template <int G>
__device__ __forceinline__ void syndrome_rows(uint32_t gmask, int d, int r, Rows& s) {
  const int up = wrap_dec(r, d);     // row i-1 (lanes >= d read a valid lane, result discarded)
  const int dn = wrap_inc(r, d);        // row i+1
  const uint32_t z0_up = gshfl<G>(gmask, s.z0, up);
  const uint32_t x1_dn = gshfl<G>(gmask, s.x1, dn);
  if (r < d) {
    s.v = s.z0 ^ z0_up ^ s.z1 ^ rotl_d(s.z1, d);
    s.pl = s.x0 ^ rotr_d(s.x0, d) ^ s.x1 ^ x1_dn;
  } else {
    s.v = 0;
    s.pl = 0;
  }
}

// Perspective selection masks (src/util.py:68-69, :77-78):
//   m0[i,j] = V[i,j] | V[i+1,j] | Pl[i,j] | Pl[i,j-1]
//   m1[i,j] = V[i,j] | V[i,j+1] | Pl[i,j] | Pl[i-1,j]
template <int G>
__device__ __forceinline__ void persp_masks(uint32_t gmask, int d, int r, uint32_t v, uint32_t pl,
                                            uint32_t& m0, uint32_t& m1) {
  const uint32_t v_dn = gshfl<G>(gmask, v, wrap_inc(r, d));
  const uint32_t pl_up = gshfl<G>(gmask, pl, wrap_dec(r, d));
  if (r < d) {
    m0 = v | v_dn | pl | rotl_d(pl, d);
    m1 = v | rotr_d(v, d) | pl | pl_up;
  } else {
    m0 = 0;
    m1 = 0;
  }
}

inline int group_size_for(int d) { return d <= 8 ? 8 : (d <= 16 ? 16 : 32); }

}  // namespace tc
