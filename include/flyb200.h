/*
 * flyb200 — C ABI of the B200-native fly_bProject hot path (libflyb200.so).
 *
 * The reference (petim0/fly_bProject) is pure Python: it has no plugin / FFI
 * of its own.  Its boundary for this path is two duck-typed classes,
 * `Fly` (reference fly.py:10-808) and `PPO` (reference ppo.py:105-279), whose
 * tensor math the entry points below replace one-for-one.  Every entry point
 * takes plain DEVICE pointers and sizes (no torch types), launches on the CUDA
 * stream passed as `stream` (a cudaStream_t cast to void*; NULL = legacy default
 * stream), never synchronises the host and allocates nothing: the caller owns
 * every buffer, exactly as the reference's in-place `self.obs_buf[:] = ...`
 * convention requires for Isaac Gym's wrapped views (fly.py:405, 418).
 *
 * Return value: 0 on success, a negative FLYB_E* code otherwise;
 * flyb_last_error() gives the message (thread-local).  There is no CPU
 * fallback: without a CUDA device every compute entry point fails.
 *
 * All floating tensors are fp32, masks/counters are int64 as in the
 * reference API (fly.py:169-179), compacted env ids are int32 as the Isaac Gym
 * `_indexed` setters take them (fly.py:457-468).
 */
#ifndef FLYB200_H_
#define FLYB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLYB_ABI_VERSION 7
#define FLYB_MAX_DOF 64      /* asset DOF count D: 18 (nmf_no_limits_limited_Dofs.urdf) or 42 (nmf_no_limits.urdf) */
#define FLYB_NUM_ABDOMEN 5   /* reference fly.py:299 */
#define FLYB_NUM_LEGTIPS 6   /* reference fly.py:300 */

#define FLYB_OK 0
#define FLYB_EINVAL (-1)     /* bad argument (null pointer, size, alignment) */
#define FLYB_ECUDA (-2)      /* CUDA runtime error (no device, launch failure) */
#define FLYB_EUNSUPPORTED (-3)

/* Constants and index tables `Fly.__init__` derives (reference fly.py:16-51, 248-348).
 * Per-env constant tensors of the reference (targets, inv_start_rot, basis vectors, origin root row,
 * initial_dofs: fly.py:121-135, 318-348) hold the same row for every env and are passed as that row. */
typedef struct FlybEnvCfg {
  int32_t num_envs;                       /* N */
  int32_t num_dof;                        /* D  = gym.get_asset_dof_count (fly.py:200) */
  int32_t num_act;                        /* A  = len(names) (fly.py:31) */
  int32_t num_bodies;                     /* B  = gym.get_asset_rigid_body_count (fly.py:201), 93 */
  int32_t action_idx[FLYB_MAX_DOF];       /* sorted DOF ids of the actuated joints, `action_indexes_one` (fly.py:288-290) */
  int32_t abdomen_local[FLYB_NUM_ABDOMEN];/* body ids of A1A2,A3..A6 inside one env, sorted (fly.py:303-314) */
  int32_t legtip_local[FLYB_NUM_LEGTIPS]; /* body ids of the six Tarsus5 links inside one env, sorted */
  float lower[FLYB_MAX_DOF];              /* dof_limits_lower (fly.py:258-267) */
  float upper[FLYB_MAX_DOF];              /* dof_limits_upper */
  float initial_pos[FLYB_MAX_DOF];        /* initial_dofs_one[:,0] (fly.py:318-336) */
  float initial_vel[FLYB_MAX_DOF];        /* initial_dofs_one[:,1] */
  float origin_root[13];                  /* origin_root_tensor row (fly.py:338-348) */
  float target[3];                        /* targets row (fly.py:134) */
  float inv_start_rot[4];                 /* fly.py:129 */
  float basis0[3];                        /* heading_vec / basis_vec0 (fly.py:127,131) */
  float basis1[3];                        /* up_vec / basis_vec1 (fly.py:125,132) */
  float dt;                               /* fly.py:16 */
  float dof_vel_scale;                    /* fly.py:43 */
  float heading_weight;                   /* fly.py:44 */
  float up_weight;                        /* fly.py:45 */
  float actions_cost_scale;               /* fly.py:46 */
  float energy_cost_scale;                /* fly.py:47 */
  float joints_at_limit_cost_scale;       /* fly.py:48 */
  float death_cost;                       /* fly.py:49 */
  float termination_height;               /* fly.py:50 */
  float termination_height_up;            /* fly.py:51 */
  float max_episode_length;               /* fly.py:34 */
  int32_t walking;                        /* 0: live standing total (fly.py:750); 1: walking total (fly.py:747-748 + leg term) */
} FlybEnvCfg;

/* Actor-critic net (reference ppo.py:10-38): trunk O->256->128, mean head 128->64->A (ELU after every layer,
 * the last one included), value head 128->64->1.  Parameters live in ONE flat fp32 device buffer so that the
 * gradient all-reduce, the global-norm clip and Adam see a single array.  Segment order is the reference's
 * `state_dict` order; the first weight matrix is stored with its rows padded from O to `k0_pad` floats and
 * every segment starts on a 16-byte boundary (padding holds zeros and receives zero gradient). */
typedef struct FlybNetLayout {
  int32_t num_obs;     /* O */
  int32_t num_act;     /* A */
  int32_t k0_pad;      /* O rounded up to a multiple of 16 */
  int32_t total;       /* floats in the flat buffer (multiple of 4) */
  int32_t off[12];     /* shared_net.0.{weight,bias}, shared_net.2.{..}, to_mean.0, to_mean.2, to_value.0, to_value.2 */
} FlybNetLayout;

/* PPO hyper-parameters (reference ppo.py:113-119, 155). */
typedef struct FlybPpoCfg {
  double clip;          /* 0.2   (doubles: the reference holds these as Python floats; torch rounds them to   */
  double lr;            /* 1e-3   fp32 only where they meet a tensor, e.g. 1 - beta1 is formed in double)     */
  double beta1, beta2;  /* 0.9, 0.999 */
  double eps;           /* 1e-8 */
  double max_grad_norm; /* 1.0 (ppo.py:198) */
} FlybPpoCfg;

/* End-of-step housekeeping of PPO.run that rides on the last launch of a step (flyb_env_post_physics) instead of
 * costing one of its own: action_var <- max(floor, action_var - decay) (reference ppo.py:236-237) and the Philox
 * draw counter += 1.  Same effect as flyb_anneal_var.  A rollout made of flyb_env_step_fused launches (which apply the
 * annealing of their step number on the fly) passes steps = T with its closing flyb_env_post_physics. */
typedef struct FlybStepTail {
  float* action_var;        /* (num_act) device; NULL = nothing to do */
  int32_t num_act;
  float decay, floor;
  uint64_t* step_counter;   /* device, nullable */
  int32_t steps;            /* how many steps' worth to apply (iterated; counter += steps); 0 counts as 1 */
} FlybStepTail;

int flyb_abi_version(void);
const char* flyb_last_error(void);
/* Number of CUDA devices visible, or a negative code. */
int flyb_device_count(void);

int flyb_net_layout(int32_t num_obs, int32_t num_act, FlybNetLayout* out);
/* Debug aid: device buffer (>= 488 uint64) that CTA 0 of the tensor-core kernels fills with clock64() marks of its
 * MMA / epilogue hand-offs; NULL (default) switches tracing off. */
int flyb_debug_set_trace(void* dev_buffer);

/* ---- tensor-core weight stream.  The GEMMs of Net.pi / Net.v (reference ppo.py:88-102) run on tcgen05 tensor
 *      cores as 3xTF32 products (fp32 accumulate; within ~1e-6 of an fp32 evaluation) when the caller passes
 *      `wpack`: the parameters re-laid as hi/lo TF32 operand slabs in the order the kernel's TMA producer
 *      consumes them.  flyb_wpack_bytes: size of that buffer, 0 when the shape is outside the tensor-core path.
 *      Two shapes are compiled: num_obs in (64, 80] with num_act <= 32 (the 18-joint actuation, O = 73: forward, act,
 *      fused env step and the PPO step on tensor cores) and anything else up to num_obs <= 160, num_act <= 48 (forward
 *      and act on tensor cores; the PPO step too where num_obs is in (144, 160] — the 42-joint set, O = 145 / A = 42;
 *      for the shapes in between flyb_ppo_grad / flyb_ppo_step run the exact-fp32 CUDA-core kernels and only Adam's
 *      in-place refresh touches the weight stream).
 *      The buffer also holds, behind the slabs, the parameter -> slab index flyb_pack_weights builds.
 *      flyb_pack_weights must be re-run whenever `params` changes by other hands (flyb_ppo_apply / flyb_ppo_step
 *      keep a given wpack in step themselves).
 *      Every entry point below that takes `wpack` accepts NULL = exact-fp32 CUDA-core path. */
size_t flyb_wpack_bytes(const FlybNetLayout* net);
int flyb_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, void* stream);

/* ---- env, kernel (a): reference Fly.get_obs + progress_buf += 1 + Fly.get_reward
 *      (fly.py:397-443, 676-681 -> compute_fly_observations 771-805, compute_fly_reward2 685-768).
 * In:  root (N,13), dof_state (N,D,2), force (N*B,3), actions_scaled (N,A) = `self.actions` (fly.py:636).
 * In/out: potentials, prev_potentials (N), progress (N) i64, reset (N) i64.
 * Out: obs (N,O), reward (N), up_vec/heading_vec (N,3), reset_ids (N) i32 = ascending ids of envs whose reset
 *      mask is now set (what `reset_buf.nonzero()` yields at fly.py:447, compacted on the device),
 *      reset_count (1) i32.  obs_copy / reward_copy (nullable) receive a second copy (rollout rows, ppo.py:228-229);
 *      reward_sum (1, nullable) += sum(reward) — the device-side accumulator behind the score log (ppo.py:233).
 *      tail (nullable): FlybStepTail applied by the last CTA to finish.
 * scratch: flyb_env_scratch_bytes(N) bytes, zeroed once by the caller before first use (a counter pair, one 64-bit
 *      reset mask per 64 envs, one tagged reset count per 128-env tile for the fused step's compaction); one scratch
 *      buffer serves flyb_env_post_physics and flyb_env_step_fused of the same env batch, on one stream. */
size_t flyb_env_scratch_bytes(int32_t num_envs);
int flyb_env_post_physics(const FlybEnvCfg* cfg, const float* root, const float* dof_state, const float* force,
                          const float* actions_scaled, float* potentials, float* prev_potentials, int64_t* progress,
                          int64_t* reset, float* obs, float* reward, float* up_vec, float* heading_vec,
                          int32_t* reset_ids, int32_t* reset_count, void* scratch, float* obs_copy,
                          float* reward_copy, float* reward_sum, const FlybStepTail* tail, void* stream);

/* ---- flyb_env_post_physics without a scratch argument.  No counterpart in the reference: it exists for hosts of this
 *      library that cannot allocate device memory — a binding that only passes the simulator's wrapped tensors through
 *      (plain ctypes / C, no torch and no CUDA-runtime binding to allocate with), or a caller inside the simulator's
 *      step loop, where cudaMalloc would synchronise the whole device or fail because the simulator took the memory up
 *      front.  The counter and reset masks the kernel needs live in an array built into the library image (~40 KB of
 *      device memory on every GPU that loads the library, held until the process exits); inputs and outputs are still
 *      the caller's.  Same arguments, same kernel and same results as flyb_env_post_physics.  The call itself
 *      allocates nothing (no cudaMalloc / cudaFree, no library routine that allocates).
 *      Limits and rules:
 *      - cfg->num_envs <= FLYB_STATIC_MAX_ENVS, else FLYB_EINVAL with a message naming the limit;
 *      - the first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
 *        current) and caches it; make that first call outside any stream capture;
 *      - all calls on a device share the one array and it keeps the previous call's contents: the entry point clears
 *        the 16-byte counter header on `stream` before every launch, so calls of different env counts may follow each
 *        other, but two calls that can overlap on the GPU (different streams without an ordering, several host threads)
 *        race on it — serialise them, or give all but one a buffer of their own through flyb_env_post_physics. */
#define FLYB_STATIC_MAX_ENVS 262144
int flyb_env_post_physics_static(const FlybEnvCfg* cfg, const float* root, const float* dof_state, const float* force,
                                 const float* actions_scaled, float* potentials, float* prev_potentials,
                                 int64_t* progress, int64_t* reset, float* obs, float* reward, float* up_vec,
                                 float* heading_vec, int32_t* reset_ids, int32_t* reset_count, float* obs_copy,
                                 float* reward_copy, float* reward_sum, const FlybStepTail* tail, void* stream);

/* ---- env: reference Fly.step up to simulate() (fly.py:626-660): scale policy actions to the joint range,
 *      scatter into the DOF position-target vector, then Fly.reset() (fly.py:446-480) for envs whose mask is set.
 * In:  actions (N,A) in [-1,1].  Out: actions_scaled (N,A), dof_targets (N*D).
 * In/out (reset): reset, progress, potentials, prev_potentials, root (N,13), dof_state (N,D,2).
 * reset == NULL: the scaling half only (the other reset-side buffers may then be NULL) — the low-gravity variant
 * resets AFTER simulate() (reference flyLowGrav.py:657-663): flyb_env_pre_physics(reset = NULL) -> simulate ->
 * flyb_env_reset -> flyb_env_post_physics. */
int flyb_env_pre_physics(const FlybEnvCfg* cfg, const float* actions, float* actions_scaled, float* dof_targets,
                         int64_t* reset, int64_t* progress, float* potentials, float* prev_potentials, float* root,
                         float* dof_state, void* stream);

/* ---- env: reference Fly.reset() on its own (fly.py:446-480; flyLowGrav.py:661-663 calls it after simulate()):
 *      envs whose mask is set get the default DOF state and the origin root row, potentials re-initialised,
 *      reset / progress cleared.  The compacted ids for the `_indexed` setters are those flyb_env_post_physics left in
 *      reset_ids / reset_count. */
int flyb_env_reset(const FlybEnvCfg* cfg, int64_t* reset, int64_t* progress, float* potentials,
                   float* prev_potentials, float* root, float* dof_state, void* stream);

/* ---- agent, kernel (b): reference PPO.run act block (ppo.py:213-220): mu = Net.pi(obs); a = mu + sqrt(var)*eps;
 *      logp = MVN(mu, diag(var)).log_prob(a); a = clip(a,-1,1) — optionally fused with flyb_env_pre_physics
 *      (pass cfg != NULL and the env buffers) so that one launch covers act + Fly.step's pre-physics half
 *      (reset == NULL with cfg != NULL: scaling + DOF targets only, no Fly.reset — low-gravity step order).
 * noise (rows,A) nullable: explicit eps (parity mode); NULL -> in-kernel Philox4x32-10 keyed by
 *   (seed, offset + *offset_dev); offset_dev (nullable) is a device counter flyb_anneal_var advances once per
 *   step, so that a captured CUDA graph draws fresh noise at every replay.
 * Out: action (rows,A) clipped, logp (rows); mu_out, obs_store (rollout all_obs row, ppo.py:226) nullable. */
int flyb_policy_act(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs, int64_t rows,
                    const float* action_var, const float* noise, uint64_t seed, uint64_t offset,
                    const uint64_t* offset_dev, float* action,
                    float* logp, float* mu_out, float* obs_store, const FlybEnvCfg* cfg, float* actions_scaled,
                    float* dof_targets, int64_t* reset, int64_t* progress, float* potentials,
                    float* prev_potentials, float* root, float* dof_state, void* stream);

/* ---- fused env step of a rollout: ONE launch = [flyb_env_post_physics of env step i-1] followed by
 *      [flyb_policy_act fused with flyb_env_pre_physics of env step i] for the same envs, i.e. reference
 *      fly.py:676-681 + ppo.py:213-220 + fly.py:626-660, 446-480 back to back.  The CTA that owns 128 envs builds their
 *      observation rows from the physics state (after simulate()) straight into the tile its tcgen05 MLP reads; the rows
 *      leave once, for the rollout ring.  Same results as the two calls, bit for bit (reward_sum up to the order of its
 *      atomics).  A T-step rollout is  flyb_policy_act (step 0), T-1 x flyb_env_step_fused, flyb_env_post_physics
 *      (closing, with FlybStepTail.steps = T): T + 1 launches instead of 2T.
 *      Needs the tensor-core path (wpack) and num_act <= 20; flyb_env_step_fused_supported() tells.
 *      Variance annealing (ppo.py:236-237): launch number i of the rollout passes anneal_steps = i — the kernel uses
 *      action_var as it will be after i annealing steps (iterated, same bits) and the Philox draw `offset + *offset_dev`
 *      with offset = i; action_var and the counter themselves are advanced by the closing post-physics tail. */
typedef struct FlybStepIO {
  /* physics state after simulate() (fly.py:378-395); root / dof_state rows of the envs this step resets are rewritten */
  float* root;               /* (N,13) */
  float* dof_state;          /* (N,D,2) */
  const float* force;        /* (N*B,3) */
  /* env buffers, in/out */
  float* actions_scaled;     /* in: `self.actions` of step i-1 (fly.py:636); out: of step i */
  float* potentials;
  float* prev_potentials;
  int64_t* progress;
  int64_t* reset;            /* out: the mask after this step's Fly.reset bookkeeping (0 for the envs it reset) */
  /* out, post-physics half */
  float* obs;                /* (N,O) nullable: env.obs_buf */
  float* obs_ring;           /* (N,O) nullable: rollout row all_obs[i] == all_next_obs[i-1] (ppo.py:226, 228) */
  float* reward;             /* (N) */
  float* reward_ring;        /* (N) nullable: all_reward[i-1] (ppo.py:229) */
  float* reward_sum;         /* (1) nullable */
  float* up_vec;             /* (N,3) */
  float* heading_vec;        /* (N,3) */
  int32_t* reset_ids;        /* (N): ascending ids of the envs whose mask step i-1 set (what step i resets) */
  int32_t* reset_count;      /* (1) */
  void* scratch;             /* flyb_env_scratch_bytes(N), zeroed once */
  /* act half */
  const float* action_var;   /* (A) */
  const float* noise;        /* (N,A) nullable: explicit eps; NULL -> Philox */
  uint64_t seed, offset;
  const uint64_t* offset_dev;
  float* action;             /* (N,A) clipped: all_acts[i] */
  float* logp;               /* (N): all_log_prob[i] */
  float* dof_targets;        /* (N*D) */
  int32_t anneal_steps;
  float anneal_decay, anneal_floor;
} FlybStepIO;
int flyb_env_step_fused_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg);
int flyb_env_step_fused(const FlybNetLayout* net, const float* params, const void* wpack, const FlybEnvCfg* cfg,
                        const FlybStepIO* io, void* stream);

/* Net.pi / Net.v forward over `rows` observation rows (ppo.py:88-102). mu_out (rows,A) / value_out (rows) nullable. */
int flyb_net_forward(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs, int64_t rows,
                     float* mu_out, float* value_out, void* stream);

/* ---- agent, kernel (c): reference PPO.make_data scan (ppo.py:160-169).
 * values (T+1,N): values[t] = v(all_obs[t]) for t<T and values[T] = v(all_next_obs[T-1]); all_next_obs[t] is
 * all_obs[t+1] for t<T-1 (ppo.py:210,228), so one value forward over T+1 rows serves both terms.
 * reward (T,N); reset_last (N) i64 = env.reset_buf after the last step: done = 1 - reset (ppo.py:230 rebinds
 * all_done to that single (N,1) mask, broadcast over T); no done mask inside the recursion.
 * Out: target (T,N), advantage (T,N). */
int flyb_gae(const float* reward, const float* values, const int64_t* reset_last, int32_t horizon, int32_t num_envs,
             float gamma, float lmbda, float* target, float* advantage, void* stream);

/* ---- data-parallel gradient exchange over NVLink peer memory (the one collective of the path).
 * One process per GPU.  Every rank allocates one exchange buffer with flyb_peer_alloc, publishes its 64-byte IPC
 * handle to the other ranks (any side channel; the Python host uses torch.distributed.all_gather_object), maps the
 * others' buffers with flyb_peer_open and fills a FlybPeerGroup.  Two forms of the exchange share the buffer.
 * FUSED (flyb_ppo_step_dp, the default of the Python host): the exchange happens inside the kernel that reduces the
 *   per-CTA partial gradients.  Every gradient entry travels as ONE 64-bit word {epoch + 1 << 32 | float bits}; 64-bit
 *   accesses are single-copy atomic, so a word that shows the awaited epoch carries its value — no fence, no flag.  The
 *   thread that has reduced four entries stores them into slot [epoch & 1][rank] of EVERY rank's buffer (posted NVLink
 *   stores), then polls the same four entries of every rank's slot in its OWN buffer and adds them in rank order
 *   (bit-identical on every rank), writes `grad` and its share of the squared norm; clip + Adam follow with no norm
 *   kernel: 4 launches, as on one GPU.  One NVLink crossing of latency.  Measured per optimiser step of 32 768 samples
 *   per rank: 138.9 us at 2 GPUs against 138.4 us on one (split form 148.5); 144.3 at 8 GPUs against 137.8 (split 151.6).
 *   A thread waits only for the peers' threads of the same index, which store before THEY wait: no dead-lock as long
 *   as the reducing grid (272 / 370 CTAs) is resident all at once on every rank — flyb_ppo_step_dp checks that against
 *   the device's occupancy and runs the SPLIT form otherwise; do not share the GPU with other long-running kernels.
 * SPLIT (flyb_ppo_grad + flyb_ppo_apply; what a caller with its own work between the halves uses):
 *   flyb_ppo_grad  : its last kernel (the fixed-order reduction of the per-CTA partial gradients) stores every
 *                    gradient entry straight into slot [epoch & 1][rank] of EVERY rank's buffer (posted NVLink
 *                    writes; every pushing CTA ends with a barrier + one system-scope fence);
 *   flyb_ppo_apply : its first kernel (global norm) raises this rank's flag on every rank, waits for all ranks'
 *                    flags of this epoch, sums the `world` slots in rank order (bit-identical on every rank), writes
 *                    `grad` and goes on as usual; the Adam kernel closes the epoch.
 *   The two calls must alternate, on one stream.  No kernel waits for a peer except that norm kernel, and what it
 *   waits for is raised at the ENTRY of the peers' norm kernels: no dead-lock however CTAs are scheduled.
 * Both: slots are double-buffered by epoch parity and a rank's next reducing kernel starts only after its current
 * exchange has ended (stream order): a rank can be at most one exchange ahead of the slowest one and never overwrites
 * what a peer still has to read.  The Adam kernel closes the epoch.  Everything is stream-ordered device code: a whole
 * data-parallel update is CUDA-graph capturable.  All ranks must use the same form for a given exchange.
 * Layout of a buffer of flyb_peer_bytes(total, world) bytes (zero-filled by flyb_peer_alloc):
 *   float slots[2][world][total_pad]; uint32 flags[FLYB_MAX_PEERS] (one epoch per source rank); uint32 epoch;
 *   uint32 error (sticky, see flyb_peer_status); time stamps (256 bytes of control in all);
 *   uint64 words[2][world][total_pad] (the fused form's slots). */
#define FLYB_MAX_PEERS 8
typedef struct FlybPeerGroup {
  int32_t rank, world;                 /* world <= FLYB_MAX_PEERS */
  int32_t total_pad;                   /* slot stride in floats: flyb_peer_total_pad(net->total) */
  int32_t timeout_ms;                  /* how long a rank waits for its peers' pushes; 0 = FLYB_PEER_TIMEOUT_S or 120 s */
  void* base[FLYB_MAX_PEERS];          /* base[r]: rank r's buffer as mapped in THIS process (base[rank]: own) */
} FlybPeerGroup;
int32_t flyb_peer_total_pad(int32_t total);
size_t flyb_peer_bytes(int32_t total, int32_t world);
/* cudaMalloc + zero + cudaIpcGetMemHandle on the current device; handle_out: 64 bytes. */
int flyb_peer_alloc(size_t bytes, void** ptr_out, void* handle_out);
int flyb_peer_open(const void* handle, void** ptr_out);   /* cudaIpcOpenMemHandle (enables peer access lazily) */
int flyb_peer_close(void* ptr);                            /* cudaIpcCloseMemHandle */
int flyb_peer_free(void* ptr);                             /* cudaFree of an own buffer */
/* Debug aid (synchronises the device): %globaltimer marks in ns of this rank's LAST exchange:
 * [0] reduce kernel entry, [3] norm kernel entry (flags raised), [4] all ranks' flags seen, [5] norm CTA 0 done;
 * fused form: [3] thread 0 has pushed its entries, [4] has seen and summed every rank's, [5] reducing CTA 0 done. */
int flyb_peer_stamps(const FlybPeerGroup* group, uint64_t* out8);
/* Sticky error word of this rank's exchange (synchronises the device): 0 = fine; 1 + r = the waiting kernel gave up
 * waiting for rank r after the timeout.  The wait does not trap (the CUDA context survives): the stream drains, every
 * later exchange returns at once, results are meaningless from then on and the host must stop — or restart with the
 * NCCL exchange (FLYB_DP=nccl), which simply waits. */
int flyb_peer_status(const FlybPeerGroup* group, int32_t* status_out);

/* ---- agent, kernel (d): one optimiser step of reference PPO.update (ppo.py:183-199), split where a
 *      data-parallel run exchanges gradients:  flyb_ppo_grad -> [sum over ranks] -> flyb_ppo_apply.
 * flyb_ppo_grad: forward, clipped-surrogate + SmoothL1 loss, backward.  wpack (nullable): packed weight stream of
 *   the CURRENT params -> tcgen05 3xTF32 forward/backward/weight-gradient GEMMs; NULL -> exact-fp32 CUDA cores.
 *   obs (rows,O), action (rows,A),
 *   old_logp/target/advantage (rows).  Out: grad (net->total) = d(loss.mean())/d(params) over THESE rows,
 *   loss (1).  workspace: flyb_ppo_workspace_bytes(net, rows) bytes.
 *   peers (nullable): also push the gradient to every rank (see FlybPeerGroup).
 * flyb_ppo_apply: grad *= grad_scale; clip_grad_norm_(max_grad_norm) with torch's coefficient
 *   min(1, max/(norm+1e-6)); Adam (torch defaults, no weight decay).  step (1) i64 device counter, incremented.
 *   grad_norm_out (1) nullable.  workspace: the flyb_ppo_grad workspace (only its first 1 KB is used).
 *   wpack (nullable): refreshed from the updated parameters by the Adam kernel itself (it must have been
 *   initialised by flyb_pack_weights, which also builds the parameter -> stream index behind the slabs).
 *   peers (nullable): grad <- sum over ranks of the pushed gradients first (pass grad_scale = 1/world);
 *   NULL: `grad` is taken as it is (single GPU, or the caller all-reduced it, e.g. with NCCL). */
size_t flyb_ppo_workspace_bytes(const FlybNetLayout* net, int64_t rows);
int flyb_ppo_grad(const FlybNetLayout* net, const FlybPpoCfg* cfg, const float* params, const void* wpack,
                  const float* obs, const float* action, const float* old_logp, const float* target,
                  const float* advantage, const float* action_var, int64_t rows, float* grad, float* loss,
                  void* workspace, const FlybPeerGroup* peers, void* stream);
int flyb_ppo_apply(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, float* grad, float* exp_avg,
                   float* exp_avg_sq, int64_t* step, float grad_scale, float* grad_norm_out, void* workspace,
                   void* wpack, const FlybPeerGroup* peers, void* stream);
/* Single-GPU optimiser step in one call: flyb_ppo_grad followed by flyb_ppo_apply(grad_scale = 1) with nothing in
 * between, which lets the library drop two launches: the gradient reduction also takes the squared norm and counts
 * the step, and Adam refreshes `wpack` in place through the index flyb_pack_weights built: 4 launches (forward/backward
 * chain, weight gradients, reduce, clip + Adam).  FLYB_PPO_TAIL=1 selects a two-launch form: the weight-gradient kernel
 * launched cooperatively, its tail — behind grid barriers — summing the per-CTA partials in the fixed order of the
 * stand-alone reduction (bit-identical gradient), taking the norm and running clip + Adam; measured 3 us slower per
 * step than the four launches, hence opt-in.
 * FLYB_ADAM_FUSED=1 selects a three-launch form (clip + Adam inside the reduction, the CTAs gathering each other's
 * tagged norm partials; bit-identical results): measured 4 us slower per step as well, opt-in.
 * Same results as the two calls up to the association of the norm's partial sums.  `grad` ends as p.grad after
 * clipping, as with flyb_ppo_apply. */
int flyb_ppo_step(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, void* wpack, const float* obs,
                  const float* action, const float* old_logp, const float* target, const float* advantage,
                  const float* action_var, int64_t rows, float* grad, float* loss, float* exp_avg, float* exp_avg_sq,
                  int64_t* step, float* grad_norm_out, void* workspace, void* stream);

/* The data-parallel optimiser step in one call: flyb_ppo_grad + gradient sum over the peer group + flyb_ppo_apply
 * (grad_scale = 1 / world) with the FUSED exchange (see FlybPeerGroup): 4 launches — chain, weight gradients, reduction
 * with the exchange inside, clip + Adam.  Same sums in the same rank order as the split form; the association of the
 * norm's partial sums is the single-GPU flyb_ppo_step's.  FLYB_DP_TAIL=1 selects the cooperative TWO-launch form instead
 * (the weight-gradient kernel reduces behind a grid barrier, pushes, raises / waits for the flags, adds the slots, takes
 * the norm and runs Adam; measured slower: 155.3 vs 146.8 us at 2 GPUs).  Without `wpack`, or for a shape the
 * tensor-core PPO step does not cover, it runs the split form. */
int flyb_ppo_step_dp(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, void* wpack, const float* obs,
                     const float* action, const float* old_logp, const float* target, const float* advantage,
                     const float* action_var, int64_t rows, float* grad, float* loss, float* exp_avg, float* exp_avg_sq,
                     int64_t* step, float grad_scale, float* grad_norm_out, void* workspace,
                     const FlybPeerGroup* peers, void* stream);

/* action_var <- max(floor, action_var - decay) (ppo.py:236-237), on the device so a captured graph can replay it;
 * step_counter (1, nullable) u64 += 1 (the Philox draw number of flyb_policy_act). */
int flyb_anneal_var(float* action_var, int32_t num_act, float decay, float floor, uint64_t* step_counter,
                    void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FLYB200_H_ */
