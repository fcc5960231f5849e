/* toric_b200.h -- C ABI of libtoric_b200.so, the B200 (sm_100a) toric-code environment engine.
 *
 * Drop-in boundary for the env hot path of Lindeby/toric-RL-decoder.  The reference has no
 * FFI of its own (it is pure Python); each entry point below names the reference interface it
 * replaces (file:line relative to the reference checkout).  INTEGRATION.md shows the ctypes
 * binding a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *     enqueued asynchronously, the library never synchronises and never allocates or frees;
 *   - d = lattice size, 3 <= d <= 32; n = number of environments;
 *   - return value: 0 ok; -1 bad argument (d out of range, null pointer, n < 0);
 *     -2 workspace / capacity too small; > 0 a cudaError_t from the launch.
 *
 * Device state layout (one lattice row per 32-bit word, bit j = column j):
 *   qubits : uint32 [n][4][d]   bit planes x0, x1, z0, z1 of the Pauli codes
 *                               (code 0:I 1:X 2:Y 3:Z -> x = code in {1,2}, z = code in {2,3})
 *   syn    : uint32 [n][2][d]   vertex plane V (= state[0]) and plaquette plane Pl (= state[1])
 */
#ifndef TORIC_B200_H
#define TORIC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TC_ABI_VERSION 6
#define TC_NUM_COUNTERS 16

/* status bits written per env */
#define TC_STATUS_BAD_ACTION 1
#define TC_STATUS_RESET_GAVE_UP 2

/* sampler modes of tc_reset / tc_actor_step */
#define TC_SAMPLE_DEPOLARIZING 0 /* generateRandomError, ToricCode.py:130-151 */
#define TC_SAMPLE_FIXED_WEIGHT 1 /* generateNRandomErrors, ToricCode.py:153-166 */
#define TC_SAMPLE_FIXED_PLUS_P 2 /* generateNPlusQRandomErrors, results/small_p_error_test.py:43-52 */
/* tc_reset only, OR-ed into the mode: one draw of the generator without reset()'s rejection of trivial
 * syndromes -- the generators called on their own (ToricCode.py:130-166) */
#define TC_SAMPLE_ONE_DRAW 16

/* flags of tc_actor_step */
#define TC_ACTOR_RANDOM_ACTIONS 1 /* draw the action on device (uniform over the perspective list x {X,Y,Z}) */
#define TC_ACTOR_AUTO_RESET 2     /* reset envs that are terminal or past max_steps (Actor_mp.py:171-175,181-183) with
                                   * the env's current p_thresh; the p_error curriculum of :176-180 is
                                   * tc_p_error_curriculum, run after the step */
#define TC_ACTOR_FREEZE_DONE 4    /* leave already-solved envs untouched (batched evaluation.py:79 loop exit) */
#define TC_ACTOR_SINGLE_PASS 8    /* tc_actor_step only: ONE kernel (step + cross-env scan by decoupled look-back +
                                   * perspective writer) instead of three; same results bit for bit.  d <=
                                   * TC_SINGLE_PASS_MAX_D; the workspace must have been zero-filled once before its
                                   * first use (it carries the look-back descriptors from launch to launch). */
#define TC_SINGLE_PASS_MAX_D 16

/* element types of tc_transitions outputs */
#define TC_DTYPE_F32 0
#define TC_DTYPE_U8 1
#define TC_DTYPE_BF16 2 /* tc_persp_write_as / tc_perspectives_as only */
#define TC_DTYPE_F16 3

int tc_abi_version(void);

/* Bytes of scratch needed by tc_perspectives / tc_actor_step for n envs of size d. */
int tc_workspace_bytes(int d, int64_t n, size_t* out_bytes);

/* Mask injection / read-back: the reference injects error masks by assigning env.qubit_matrix and
 * env.state (results/small_p_error_test.py:119-120).  codes: uint8 [n][2][d][d] Pauli codes. */
int tc_pack_qubits(const uint8_t* codes, uint32_t* qubits, int d, int64_t n, void* stream);
int tc_unpack_qubits(const uint32_t* qubits, uint8_t* codes, int d, int64_t n, void* stream);
/* state01: uint8 [n][2][d][d] of 0/1 (any non-zero counts as a defect when packing). */
int tc_pack_state(const uint8_t* state01, uint32_t* syn, int d, int64_t n, void* stream);
int tc_unpack_state(const uint32_t* syn, uint8_t* state01, int d, int64_t n, void* stream);
/* The same state as lattice-row bitmasks: rows uint8 [n][2][d][ceil(d/8)] = numpy
 * packbits(state, axis=-1, bitorder="little") -- the compact host format of the new states (a step result of
 * EnvSet.step, src/EnvSet.py:38-47) for callers that read every step's states back over PCIe. */
int tc_state_rows(const uint32_t* syn, uint8_t* rows, int d, int64_t n, void* stream);

/* createSyndromOpt (ToricCode.py:168-199): syn <- stencil(qubits). */
int tc_syndrome(const uint32_t* qubits, uint32_t* syn, int d, int64_t n, void* stream);

/* ToricCode.step + getReward + isTerminalState (ToricCode.py:62-88, :202-218), batched like
 * EnvSet.step (src/EnvSet.py:38-47).  actions: int32 [n][4] = (matrix, row, col, op).  The syndrome
 * is updated incrementally.  reward: float32 [n] (100 on terminal else defect difference);
 * terminal: uint8 [n]; status: uint8 [n] (OR-ed with TC_STATUS_BAD_ACTION), nullable. */
int tc_step(uint32_t* qubits, uint32_t* syn, const int32_t* actions, float* reward, uint8_t* terminal,
            uint8_t* status, int d, int64_t n, void* stream);

/* ToricCode.reset (ToricCode.py:90-127) for the envs listed in idx (int64 [n_idx], NULL = envs
 * 0..n_idx-1), like EnvSet.resetAll / resetTerminalEnvs (src/EnvSet.py:19-36).  Sampling is
 * Philox4x32-10 keyed by `seed` with counter (env_base + env, episode[env], draw); episode[env] is
 * incremented.  p_thresh: uint32 [n] per-env error threshold floor(p * 2^32).  gs (sticky
 * ground-state flag) is set to 1.  n_fixed = number of forced errors for the fixed-weight modes. */
int tc_reset(uint32_t* qubits, uint32_t* syn, uint8_t* gs, uint8_t* status, const uint32_t* p_thresh,
             uint32_t* episode, const int64_t* idx, int64_t n_idx, int mode, int n_fixed, uint64_t seed,
             int64_t env_base, int d, void* stream);

/* evalGroundState (ToricCode.py:289-337): gs[e] &= all four plane parities even. */
int tc_ground_state(const uint32_t* qubits, uint8_t* gs, int d, int64_t n, void* stream);

/* generatePerspectiveBatch + the concatenate / float32 layout of selectActionBatch
 * (src/numba/util_actor.py:56-67 and :33-39; single-env forms src/util.py:46-150).
 *   out      float32 [cap][2][d][d]  perspectives, env-major, per env all matrix-0 qubits row-major
 *                                    then all matrix-1 qubits (rotated views)
 *   pos      uint8   [cap][4]        (m, i, j, 0) per perspective                     (nullable)
 *   counts   int32   [n]             perspectives per env
 *   offsets  int64   [n+1]           exclusive prefix of counts; offsets[n] = total (= splice_idx)
 * Envs whose perspectives would not fit in cap are skipped and *overflow (int32, nullable) is set.
 * workspace: tc_workspace_bytes(d, n) bytes. */
int tc_perspectives(const uint32_t* syn, float* out, uint8_t* pos, int32_t* counts, int64_t* offsets,
                    int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes, int d,
                    int64_t n, void* stream);

/* tc_perspectives without the workspace argument -- NO COUNTERPART IN THE REFERENCE (which is Python and allocates
 * freely); it exists for the caller's sake.  Same kernels, same outputs bit for bit.
 *
 * Who needs it: code that hands this library device buffers it was itself given and that has no way to get more
 * device memory at the point of the call -- an operator inside a host framework whose memory pool it may not touch
 * (cudaMalloc from there bypasses the pool, and cudaFree synchronises the whole device), a process that gave all
 * device memory to one arena at start-up so that the driver has nothing left to hand out, or a real-time decoding
 * loop that must not wait on the allocator's lock.  Such a caller still owns its inputs and
 * outputs (syn, out, pos, counts, offsets, overflow) but cannot produce the tc_workspace_bytes(d, n) scratch
 * block, so the scratch comes from an array compiled into the library instead.
 *
 * The price and the limit: the array is TC_STATIC_SCRATCH_BYTES (1 MiB) of device memory on EVERY GPU that loads
 * the library, held until the process exits even if this entry point is never called.  A batch is accepted when
 * tc_workspace_bytes(d, n) <= TC_STATIC_SCRATCH_BYTES -- at least TC_STATIC_MAX_ENVS (65 536) envs for every d;
 * anything larger returns -2 and prints one line to stderr saying how many bytes were needed (split the batch, or
 * use tc_perspectives).  The library calls no cudaMalloc / cudaFree / allocating routine on this path.
 *
 * First call on a device: the library looks up that device's address of the array (cudaGetSymbolAddress, with the
 * device of the call current) and sets the writer kernel's shared-memory attribute.  Make that first call per
 * device during start-up, before any stream capture begins.  Later calls only enqueue kernels.  The array used is
 * the CURRENT device's: syn and the outputs must live there, as for every other entry point.
 *
 * Sharing: there is ONE array per device for all callers, streams and threads of the process, and it keeps
 * whatever the last call left in it.  That is harmless from call to call (every word this path reads is rewritten
 * by the same call first, nothing needs clearing), but two calls whose kernels can run at the same time on one GPU
 * -- different streams, or different host threads -- overwrite each other's prefix sums and produce garbage without
 * any error.  Serialise them: one stream, or an event between the calls. */
#define TC_STATIC_SCRATCH_BYTES (1u << 20)
#define TC_STATIC_MAX_ENVS 65536
int tc_perspectives_static(const uint32_t* syn, float* out, uint8_t* pos, int32_t* counts, int64_t* offsets,
                           int32_t* overflow, int64_t cap, int d, int64_t n, void* stream);

/* The two halves of tc_perspectives, for callers that time or overlap them separately:
 * count + scan (fills counts, offsets[n] and the workspace), then the streaming write. */
int tc_persp_count(const uint32_t* syn, int32_t* counts, int64_t* offsets, void* workspace,
                   size_t workspace_bytes, int d, int64_t n, void* stream);
int tc_persp_write(const uint32_t* syn, float* out, uint8_t* pos, int64_t* offsets, int32_t* overflow,
                   int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n, void* stream);

/* tc_persp_write / tc_perspectives with the element type of the network-input tensor chosen by the caller
 * ("T* out, int out_dtype").  The reference casts to float32
 * (`.type('torch.Tensor')`, src/numba/util_actor.py:39; src/util_actor.py:36); the values are 0/1, so
 * TC_DTYPE_BF16 / TC_DTYPE_F16 / TC_DTYPE_U8 carry the same tensor exactly at 2 / 2 / 1 bytes per element
 * for a network run under bf16/fp16 autocast or with an integer input stage.  out: [cap][2][d][d] of that
 * type, 16-byte aligned. */
int tc_persp_write_as(const uint32_t* syn, void* out, int out_dtype, uint8_t* pos, int64_t* offsets,
                      int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes, int d, int64_t n,
                      void* stream);
int tc_perspectives_as(const uint32_t* syn, void* out, int out_dtype, uint8_t* pos, int32_t* counts,
                       int64_t* offsets, int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes,
                       int d, int64_t n, void* stream);

/* One fused actor iteration over all envs (the loop body of src/Actor_mp.py:104-185 minus the
 * network): optional on-device random action, step, reward/terminal, episode bookkeeping with
 * auto-reset, then the perspectives of the resulting states.
 *   actions     int32 [n][4]  read (host policy) or written (TC_ACTOR_RANDOM_ACTIONS)
 *   steps       int32 [n]     steps taken in the current episode
 *   counters    int64 [TC_NUM_COUNTERS] running totals, nullable
 *   prev_syn    uint32 [n][2][d] copy of the pre-reset post-step syndrome (transition next_state), nullable
 *   gs          uint8 [n]     out, a per-step RESULT like reward / terminal: 1 iff all four plane parities are
 *                             even after this step and before any auto-reset, i.e. what evalGroundState
 *                             (ToricCode.py:326-335) returns on the post-step state with a fresh flag -- for an
 *                             episode that ends in this step it is the episode's logical outcome.  It is
 *                             OVERWRITTEN every step (not AND-ed) and not touched for frozen envs, so it is not
 *                             the sticky flag of tc_ground_state / tc_reset: pass a different buffer when both
 *                             interfaces are used on the same envs (ToricEngine: `step_gs` vs `gs`).
 */
int tc_actor_step(uint32_t* qubits, uint32_t* syn, uint32_t* prev_syn, int32_t* actions, float* reward,
                  uint8_t* terminal, uint8_t* gs, uint8_t* status, int32_t* steps, uint32_t* episode,
                  const uint32_t* p_thresh, int64_t* counters, int flags, int max_steps, int mode,
                  int n_fixed, uint64_t seed, int64_t env_base, float* out, uint8_t* pos, int32_t* counts,
                  int64_t* offsets, int32_t* overflow, int64_t cap, void* workspace, size_t workspace_bytes,
                  int d, int64_t n, void* stream);

/* tc_actor_step without the perspective write (follow with tc_persp_write on the same workspace). */
int tc_actor_pre(uint32_t* qubits, uint32_t* syn, uint32_t* prev_syn, int32_t* actions, float* reward,
                 uint8_t* terminal, uint8_t* gs, uint8_t* status, int32_t* steps, uint32_t* episode,
                 const uint32_t* p_thresh, int64_t* counters, int flags, int max_steps, int mode, int n_fixed,
                 uint64_t seed, int64_t env_base, int32_t* counts, int64_t* offsets, void* workspace,
                 size_t workspace_bytes, int d, int64_t n, void* stream);

/* Host-buffer form of one actor iteration: what the reference's actor does around EnvSet.step with HOST arrays
 * (src/Actor_mp.py:116-125, src/EnvSet.py:38-47) -- actions come from pinned host memory, rewards / terminals /
 * perspective counts / new states go back to pinned host memory, the perspective tensor stays on the device
 * (`.to(device)`, src/numba/util_actor.py:39).  ONE call enqueues, without ever blocking the host,
 *   h2d stream:      actions -> d_actions;                                  record ev_actions
 *   compute stream:  wait ev_actions [, ev_unpacked_prev]; tc_actor_pre (results written straight into d_stage);
 *                    record ev_compute; tc_persp_write_as
 *   d2h stream:      wait ev_compute; state_mode 1: tc_unpack_state -> d_stage + o_state (state_mode 2, the
 *                    tc_state_rows format, is written by the step kernel itself; 0: no states); record ev_unpacked;
 *                    d_stage -> h_stage; record ev_result
 * so the step's results, final before the perspective writer runs, cross PCIe underneath the writer (give the d2h
 * stream a higher priority than the compute stream: its work must not queue behind the writer's CTAs).  The caller
 * waits on ev_result (cudaEventSynchronize) before it reads h_stage.  Streams and events are the caller's
 * (cudaStream_t / cudaEvent_t as void*); the library creates nothing.  d_stage / h_stage layout: float32 reward [n]
 * at o_reward, int32 counts [n] at o_counts, uint8 terminal [n] at o_term, states at o_state (16-byte aligned).
 * ev_unpacked_prev (nullable): the ev_unpacked of the previous call -- this step's kernels rewrite the syndromes
 * that call's state kernel reads. */
typedef struct tc_host_step {
  uint32_t* qubits; uint32_t* syn; uint32_t* prev_syn; uint8_t* gs; uint8_t* status; int32_t* steps;
  uint32_t* episode; const uint32_t* p_thresh; int64_t* counters; int64_t* offsets; int32_t* overflow;
  void* out; uint8_t* pos; int64_t cap; int out_dtype;
  void* workspace; size_t workspace_bytes;
  int flags, max_steps, mode, n_fixed; uint64_t seed; int64_t env_base; int d; int64_t n;
  const int32_t* h_actions; int32_t* d_actions; uint8_t* d_stage; uint8_t* h_stage; size_t stage_bytes;
  size_t o_reward, o_counts, o_term, o_state; int state_mode;
  void* h2d_stream; void* compute_stream; void* d2h_stream;
  void* ev_actions; void* ev_compute; void* ev_unpacked_prev; void* ev_unpacked; void* ev_result;
} tc_host_step;
int tc_host_actor_step(const tc_host_step* s);

/* ---- policy-side entry points (the callers either side of the env path) ---- */

/* _selectActionBatch_prime + the epsilon-greedy draw of selectActionBatch (src/numba/util_actor.py:41-107;
 * single-env form src/util_actor.py:47-65).  One segment of Q-value rows per env:
 *   q_values  float32 [sumP][3]   network output for the perspectives of tc_perspectives / tc_actor_step
 *   pos       uint8   [sumP][4]   their (m,i,j,0) positions;  offsets int64 [n+1]
 *   actions   int32   [n][4]      out: (m, i, j, a+1) of the chosen (perspective, action)
 *   q_sel     float32 [n][3]      out: the Q-value row of the chosen perspective (reference's q_v)
 *   sel_idx   int64   [n]         out: global row of the chosen perspective, -1 when the env has none (nullable)
 *   was_greedy uint8  [n]         out: 1 greedy / 0 random (nullable)
 * Greedy = first maximum in row-major (perspective, action) order (np.argmax tie-break).  The decision is
 * drawn on device: Philox(seed; env_base+e, episode[e], 0x60000000|(steps[e] & 0x1fffffff)) -> random iff u0 < eps
 * threshold (eps_thresh uint32 [n] per env, or eps_scalar when NULL; threshold = floor(eps * 2^32)),
 * random perspective mulhi(u1, P), random action mulhi(u2, 3); greedy_in (uint8 [n], nullable) overrides
 * the decision.  episode / steps nullable (0).  Envs without perspectives get the identity action
 * (0,0,0,0) and q_sel = 0. */
int tc_select_action(const float* q_values, const uint8_t* pos, const int64_t* offsets, const uint32_t* episode,
                     const int32_t* steps, const uint32_t* eps_thresh, uint32_t eps_scalar,
                     const uint8_t* greedy_in, uint64_t seed, int64_t env_base, int32_t* actions, float* q_sel,
                     int64_t* sel_idx, uint8_t* was_greedy, int64_t n, void* stream);

/* The actor's p_error curriculum (src/Actor_mp.py:43-47,171-181; Distributed_mp.py:53-56,186-189) without leaving the
 * device.  All probabilities are thresholds floor(p * 2^32).  For every env that was auto-reset by the fused step just
 * before (steps[e] == 0; all_envs != 0: every env -- call once that way after the initial reset to prime):
 *   p_cap[e]    <- min(final, p_cap[e] + delta)                                   (:176; p_cap starts at `start`, :47)
 *   p_thresh[e] <- p_cap[e]                                       'linear'        (:180)
 *               or uniform on the integers [min(start, cap), cap] 'random'        (:178), drawn as
 *                  lo + mulhi(Philox(seed; env_base+e, episode[e], 0xA0000000).x, cap - lo + 1)
 * p_thresh[e] is what the env's NEXT auto-reset samples with: its k-th reset sees the reference's k-th ceiling, chosen
 * one episode early, so the fused step itself does not change.  steps int32 [n], episode uint32 [n] (read);
 * p_cap, p_thresh uint32 [n] (read and written). */
int tc_p_error_curriculum(const int32_t* steps, const uint32_t* episode, uint32_t* p_cap, uint32_t* p_thresh,
                           uint32_t start_thresh, uint32_t delta_thresh, uint32_t final_thresh, int random_strategy,
                           int all_envs, uint64_t seed, int64_t env_base, int64_t n, void* stream);

/* generateTransitionParallel's state pair (src/util_actor.py:223-264): both syndromes rolled so the acted
 * qubit sits at (g,g) (shift_state, src/util.py:97-102) and rotated when it lives in matrix 1
 * (rotate_state, src/util.py:87-94).  out_prev / out_next: [n][2][d][d] of out_dtype (TC_DTYPE_F32 or
 * TC_DTYPE_U8), either nullable;  action_out int32 [n][4] = (m, g, g, op), nullable. */
int tc_transitions(const uint32_t* syn_prev, const uint32_t* syn_next, const int32_t* actions, void* out_prev,
                   void* out_next, int32_t* action_out, int out_dtype, int d, int64_t n, void* stream);

/* computePrioritiesParallel (src/util_actor.py:268-287): |reward + discount * max_a q_next[a] - q[op-1]|.
 * actions int32 [n][4]; q, q_next float32 [n][3]; priorities float32 [n]. */
int tc_td_priorities(const int32_t* actions, const float* reward, const float* q, const float* q_next,
                     float discount, float* priorities, int64_t n, void* stream);

/* The per-state maximum of predictMaxOptimized (src/util_learner.py:88-111): max over the state's Q-value
 * segment; states with fewer perspectives than pad_to (the batch maximum; read from *pad_to_dev when that
 * is non-NULL) also compete against 0 (the reference zero-pads, :100-106); states without perspectives
 * give 0 (:109). */
int tc_segmented_max(const float* q_values, const int64_t* offsets, int pad_to, const int32_t* pad_to_dev,
                     float* out, int64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TORIC_B200_H */
