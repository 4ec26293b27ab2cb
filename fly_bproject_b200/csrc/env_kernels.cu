// Kernel (a): fused observation + reward + termination + reset-id compaction, and the pre-physics half of
// Fly.step (action scaling, DOF-target scatter, reset bookkeeping).
//
// Replaces reference fly.py:397-443, 446-480, 624-660, 676-681 and the two TorchScript functions
// compute_fly_observations (771-805) / compute_fly_reward2 (685-768): ~385 aten launches per step in the
// reference become one launch each side of the physics step.
//
// Data movement (HBM-bound; algorithmic bytes per env-step in DESIGN.md):
//   * one CTA owns 64 consecutive envs, four threads per env; the root slab comes in as one coalesced float4
//     sweep, obs rows are assembled in shared memory and leave as full 128-bit stores (rows are 52 / 292 bytes,
//     so per-thread row access would be misaligned);
//   * the 11 contact-force rows an env needs (5 abdomen + 6 leg tips out of 93 bodies) and its actuated DOF
//     pairs are gathered by the env's four threads with stride 4 through shared-memory offset tables;
//   * reset ids are compacted in ascending order (== torch.nonzero) inside the same launch: warp ballot ->
//     one 64-bit mask per CTA -> the last CTA to finish scans the bitmap and writes the id list; the count
//     stays on the device — the reference's host sync at fly.py:447-450 is gone.
#include <cstdlib>
#include <mutex>

#include "env_math.cuh"
#include "env_scan.cuh"
#include "tc_common.cuh"

namespace flyb {
unsigned long long* tc_get_trace();

constexpr int ENVS = 64;           // envs per CTA
constexpr int PP_THREADS = 256;    // 4 threads per env for the load / store sweeps; threads [0, ENVS) own one env each
constexpr int PP_WARPS = PP_THREADS / 32;

// The kernel is latency-bound at the env counts of the reference's configs (16384 envs = 12.5 MB per launch), so
// every thread issues ALL its global loads before the first dependent shared-memory store, with as little address
// arithmetic in front of them as possible: a thread is (env e = tid / 4, quarter qd = tid % 4) and walks its env's
// 33 gathered contact values / A actuated DOFs with stride 4 through small shared-memory offset tables.
// Reset ids: every CTA publishes a 64-bit mask of its envs; the LAST CTA to finish (a counter tells) scans the
// bitmap and writes the ascending id list (== torch.nonzero) and the count — no CTA ever waits for another one.
__global__ void __launch_bounds__(PP_THREADS)
post_physics_kernel(const __grid_constant__ FlybEnvCfg cfg, const float* __restrict__ root,
                    const float* __restrict__ dof_state, const float* __restrict__ force,
                    const float* __restrict__ actions_scaled, float* __restrict__ potentials,
                    float* __restrict__ prev_potentials, long long* __restrict__ progress,
                    long long* __restrict__ reset, float* __restrict__ obs, float* __restrict__ reward,
                    float* __restrict__ up_vec, float* __restrict__ heading_vec, int* __restrict__ reset_ids,
                    int* __restrict__ reset_count, ScanScratch* __restrict__ scratch, float* __restrict__ obs_copy,
                    float* __restrict__ reward_copy, float* __restrict__ reward_sum,
                    const __grid_constant__ FlybStepTail tail, unsigned long long* __restrict__ trace,
                    const int force_bulk) {
  // debug (flyb_debug_set_trace): thread 0 of CTA 0 marks [160..]: entry, loads issued, loads consumed, per-env math
  // done, reward + masks done, mask published, obs slab out (clock64 is per SM: marks of other CTAs would not compare)
  int tmark = 160;
  auto mark = [&]() {
    if (trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && tmark < 170) trace[tmark++] = clock64();
  };
  mark();
  extern __shared__ __align__(16) float smem[];
  const int N = cfg.num_envs, D = cfg.num_dof, A = cfg.num_act, B = cfg.num_bodies;
  const int O = 19 + 3 * A;
  float* obs_s = smem;                        // [ENVS][O]      (O is odd -> conflict-free row access)
  float* root_s = obs_s + ENVS * O;           // [ENVS][13]
  float* force_s = root_s + ENVS * 13;        // [ENVS][33]
  // force_bulk: the CTA's 64 contiguous contact-force blocks (64 x B x 12 bytes) arrive as ONE bulk copy and the 33
  // values an env needs are picked out of shared memory.  Gathering 11 of 93 rows through the LSU touches ~940 of an
  // env's 1116 bytes in DRAM anyway (64-byte access granularity) and runs at the SM's outstanding-miss limit
  // (~1.5 TB/s machine-wide, measured); the sequential copy streams at the HBM rate.
  float* force_blk = force_s + ENVS * FORCE_VALS + (ENVS * (O + 13 + FORCE_VALS) & 3 ? 4 - (ENVS * (O + 13 + FORCE_VALS) & 3) : 0);
  __shared__ __align__(8) uint64_t s_fbar;
  __shared__ unsigned s_mask[ENVS / 32];
  __shared__ int s_cnt[ENVS];                 // role 3 -> role 0: legs | at_limit << 8 | abdomen_hit << 16
  __shared__ float s_sq[ENVS];
  __shared__ int s_scan[PP_WARPS];
  __shared__ int s_last;
  // small tables: per-lane indexed reads of the __grid_constant__ copy would serialise in the constant cache
  __shared__ int s_foff[FORCE_VALS];          // float offset of gathered value rc inside one env's force block
  __shared__ int s_aidx[FLYB_MAX_DOF];
  __shared__ float s_lo[FLYB_MAX_DOF], s_hi[FLYB_MAX_DOF];

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int blk = blockIdx.x;
  const int env0 = blk * ENVS;
  const int nenv = min(ENVS, N - env0);
  static_assert(PP_THREADS == 4 * ENVS, "four threads per env");
  const int e = tid >> 2, qd = tid & 3;
  const bool e_ok = e < nenv;
  const unsigned fbytes = (unsigned)nenv * (unsigned)B * 12u;
  const bool bulk = force_bulk != 0 && (fbytes & 15u) == 0u;   // 64 x 1116 bytes per full CTA: a multiple of 16
  if (bulk && tid == 0) {
    tc::mbar_init(&s_fbar, 1);
    tc::fence_mbar_init();
    tc::mbar_arrive_expect_tx(&s_fbar, fbytes);
    tc::tma_load_1d(force_blk, reinterpret_cast<const char*>(force) + (size_t)env0 * B * 12, fbytes, &s_fbar);
  }

  // ---- loads, wave 1: root slab (contiguous nenv*13 floats; 16B aligned since ENVS*52 % 16 == 0), per-env scalars
  constexpr int ROOT4 = ENVS * 13 / 4;        // 208 <= PP_THREADS
  static_assert(ROOT4 <= PP_THREADS, "one float4 of the root slab per thread");
  const int nroot = nenv * 13;
  float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (tid * 4 + 3 < nroot) r4 = ldg_stream4(reinterpret_cast<const float4*>(root + (size_t)env0 * 13) + tid);
  float pot_in = 0.0f;
  long long prog_in = 0, reset_in = 0;
  if (tid < nenv) {
    pot_in = potentials[env0 + tid];
    prog_in = progress[env0 + tid];
    reset_in = reset[env0 + tid];
  }
  if (tid < FORCE_VALS) {
    const int r = tid / 3;
    const int body = (r < FLYB_NUM_ABDOMEN) ? cfg.abdomen_local[r] : cfg.legtip_local[r - FLYB_NUM_ABDOMEN];
    s_foff[tid] = body * 3 + (tid - r * 3);
  }
  if (tid >= 64 && tid < 64 + A) {
    const int d = cfg.action_idx[tid - 64];
    s_aidx[tid - 64] = d;
    s_lo[tid - 64] = cfg.lower[d];
    s_hi[tid - 64] = cfg.upper[d];
  }
  __syncthreads();
  // ---- wave 2: the 11 gathered contact-force rows of this thread's env, values rc = qd, qd + 4, ...
  constexpr int FITER = (FORCE_VALS + 3) / 4;   // 9
  float fv[FITER];
  if (!bulk) {
    const float* fe = force + (size_t)(env0 + e) * B * 3;
#pragma unroll
    for (int it = 0; it < FITER; ++it) {
      const int rc = qd + 4 * it;
      fv[it] = (e_ok && rc < FORCE_VALS) ? __ldg(fe + s_foff[rc]) : 0.0f;
    }
  }
  // ---- wave 3: actuated DOF (pos, vel) pairs and commanded actions of this thread's env, k = qd, qd + 4, ...
  //      -> obs[12 : 12+A], obs[12+A : 12+2A], obs[12+2A : 12+3A]   (fly.py:795, 801, 802)
  {
    const float2* de = reinterpret_cast<const float2*>(dof_state) + (size_t)(env0 + e) * D;
    const float* ae = actions_scaled + (size_t)(env0 + e) * A;
    float* oe = obs_s + e * O + 12;
    constexpr int DU = 5;                        // 20 DOFs per pass: one pass for the 18-DOF asset
    for (int k0 = 0; k0 < A; k0 += 4 * DU) {
      float2 pv[DU];
      float act[DU];
#pragma unroll
      for (int u = 0; u < DU; ++u) {
        const int k = k0 + qd + 4 * u;
        pv[u] = make_float2(0.f, 0.f);
        act[u] = 0.0f;
        if (e_ok && k < A) {
          pv[u] = __ldg(de + s_aidx[k]);
          act[u] = __ldg(ae + k);
        }
      }
#pragma unroll
      for (int u = 0; u < DU; ++u) {
        const int k = k0 + qd + 4 * u;
        if (e_ok && k < A) {
          oe[k] = unscale(pv[u].x, s_lo[k], s_hi[k]);
          oe[A + k] = pv[u].y * cfg.dof_vel_scale;
          oe[2 * A + k] = act[u];
        }
      }
    }
  }
  mark();   // every load of this thread has been issued; the first dependent use follows
  if (bulk) {
    tc::mbar_wait(&s_fbar, 0u);
    const float* fe = force_blk + (size_t)e * B * 3;   // block stride B * 3 = 279 words: odd, conflict-free across envs
#pragma unroll
    for (int it = 0; it < FITER; ++it) {
      const int rc = qd + 4 * it;
      fv[it] = (e_ok && rc < FORCE_VALS) ? fe[s_foff[rc]] : 0.0f;
    }
  }
  if (tid * 4 + 3 < nroot) reinterpret_cast<float4*>(root_s)[tid] = r4;
  for (int i = (nroot & ~3) + tid; i < nroot; i += PP_THREADS) root_s[i] = __ldg(root + (size_t)env0 * 13 + i);
#pragma unroll
  for (int it = 0; it < FITER; ++it) {
    const int rc = qd + 4 * it;
    if (e_ok && rc < FORCE_VALS) force_s[e * FORCE_VALS + rc] = fv[it];
  }
  __syncthreads();
  mark();

  // ---- per-env math, split over four warp pairs (independent dependency chains instead of one long one; a role is
  //      a whole warp, so nothing diverges):  role 0: potential, heading / up projections, later reward + termination;
  //      role 1: local velocities;  role 2: Euler angles, angle to target;  role 3: contact flags, action / limit sums.
  static_assert(PP_WARPS == 8 && ENVS == 64, "four roles x two warps x 32 envs");
  const int role = warp >> 1;
  const int me = ((warp & 1) << 5) | lane;      // env of this thread inside the CTA (role 0: me == tid)
  const bool me_ok = me < nenv;
  EnvObsHead h;
  h.heading_proj = 0.0f; h.potential = 0.0f;
  float r[13];
#pragma unroll
  for (int j = 0; j < 13; ++j) r[j] = me_ok ? root_s[me * 13 + j] : (j == 6 ? 1.0f : 0.0f);
  {
    float* o = obs_s + me * O;
    const Quat tq = env_torso_quat(r, cfg);
    if (role == 0) {
      env_head_heading(r, cfg, tq, h);
      if (me_ok) { o[10] = h.up_proj; o[11] = h.heading_proj; }
    } else if (role == 1) {
      env_head_vel(r, tq, h);
      if (me_ok) {
        o[0] = h.height;
        o[1] = h.vel_loc.x; o[2] = h.vel_loc.y; o[3] = h.vel_loc.z;
        o[4] = h.angvel_loc.x; o[5] = h.angvel_loc.y; o[6] = h.angvel_loc.z;
      }
    } else if (role == 2) {
      env_head_euler(r, cfg, tq, h);
      if (me_ok) { o[7] = h.yaw; o[8] = h.roll; o[9] = h.angle_to_target; o[12 + 3 * A] = h.pitch; }
    } else if (me_ok) {
      const float* f = force_s + me * FORCE_VALS;
      float abdomen = 0.0f;
#pragma unroll
      for (int b = 0; b < FLYB_NUM_ABDOMEN; ++b) abdomen += (f[b * 3] + f[b * 3 + 1]) + f[b * 3 + 2];
      int legs = 0, at_limit = 0;
#pragma unroll
      for (int b = 0; b < FLYB_NUM_LEGTIPS; ++b) {
        const float* fr = f + (FLYB_NUM_ABDOMEN + b) * 3;
        const bool touch = ((fr[0] + fr[1]) + fr[2]) > 0.0f;
        o[13 + 3 * A + b] = touch ? 1.0f : 0.0f;  // fly.py:797, 803
        legs += touch ? 1 : 0;
      }
      float actions_sq_sum = 0.0f;
      const float* cmd = o + 12 + 2 * A;
      for (int k = 0; k < A; ++k) {
        const float av = cmd[k];
        actions_sq_sum += av * av;                 // fly.py:731 (the electricity term of fly.py:733 is identically 0)
        at_limit += (av > s_hi[k] * 0.9f) ? 1 : 0;  // fly.py:736-737
        at_limit += (av < s_lo[k] * 0.9f) ? 1 : 0;
      }
      s_cnt[me] = legs | (at_limit << 8) | ((abdomen > 0.0f ? 1 : 0) << 16);
      s_sq[me] = actions_sq_sum;
    }
  }
  __syncthreads();
  mark();
  bool flag = false;
  float my_reward = 0.0f;
  if (role == 0 && me_ok) {
    const int env = env0 + me;
    const int packed = s_cnt[me];
    EnvRewardIn in;
    in.actions_sq_sum = s_sq[me];
    in.electricity = 0.0f;
    in.at_limit = (packed >> 8) & 0xff;
    in.height = r[2];
    in.heading_proj = h.heading_proj;
    in.qz = r[5];
    in.qw = r[6];
    in.prev_potential = pot_in;  // fly.py:780
    in.potential = h.potential;
    in.legs_touching = packed & 0xff;
    in.abdomen_hit = (packed >> 16) != 0;
    long long prog = prog_in + 1;  // fly.py:678
    if (prog == 0) prog = 1;       // fly.py:415-416
    in.progress = prog;
    in.reset_prev = reset_in;
    float rew;
    long long rst;
    env_reward_reset(in, cfg, &rew, &rst);
    potentials[env] = h.potential;
    prev_potentials[env] = pot_in;
    progress[env] = prog;
    reset[env] = rst;
    reward[env] = rew;
    if (reward_copy) reward_copy[env] = rew;
    my_reward = rew;
    up_vec[env * 3 + 0] = h.up_vec.x; up_vec[env * 3 + 1] = h.up_vec.y; up_vec[env * 3 + 2] = h.up_vec.z;
    heading_vec[env * 3 + 0] = h.heading_vec.x; heading_vec[env * 3 + 1] = h.heading_vec.y;
    heading_vec[env * 3 + 2] = h.heading_vec.z;
    flag = rst != 0;
  }
  if (role == 0) {
    const unsigned ballot = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_mask[warp] = ballot;
    if (reward_sum) {   // score accumulator (ppo.py:233), one atomic per warp
      const float ws = warp_sum(my_reward);
      if (lane == 0) atomicAdd(reward_sum, ws);
    }
  }
  __syncthreads();
  mark();
  static_assert(ENVS == 64, "one 64-bit reset mask per CTA");
  unsigned long long* bitmap = scratch->status;
  // publish the mask and count this CTA as done BEFORE its obs slab leaves: the CTA that sees the counter complete
  // only needs the masks, so its scan overlaps the other CTAs' stores (the atomic's result is consumed after the slab)
  unsigned done_prev = 0u;
  if (tid == 0) {
    st_status(bitmap + blk, ((unsigned long long)s_mask[1] << 32) | (unsigned long long)s_mask[0]);
    __threadfence();
    done_prev = atomicAdd(&scratch->done, 1u);
  }
  mark();

  // ---- obs slab out: full 128-bit stores (slab base 16B aligned: ENVS*O*4 % 16 == 0)
  {
    const int n = nenv * O;
    const int n4 = n >> 2;
    float4* dst4 = reinterpret_cast<float4*>(obs + (size_t)env0 * O);
    // the rollout ring row (obs_copy) starts at (i + 1) * N * O floats: 16-byte aligned only when N * O is a multiple
    // of 4 (N = 1001, O = 73 is not) -> that copy then leaves as scalar stores
    const bool cpy_vec = obs_copy != nullptr && (reinterpret_cast<uintptr_t>(obs_copy) & 15u) == 0;
    float4* cpy4 = cpy_vec ? reinterpret_cast<float4*>(obs_copy + (size_t)env0 * O) : nullptr;
    for (int i = tid; i < n4; i += PP_THREADS) {
      const float4 v = reinterpret_cast<const float4*>(obs_s)[i];
      dst4[i] = v;
      if (cpy4) stg_stream4(cpy4 + i, v);
    }
    for (int i = (n4 << 2) + tid; i < n; i += PP_THREADS) {
      obs[(size_t)env0 * O + i] = obs_s[i];
      if (cpy_vec) obs_copy[(size_t)env0 * O + i] = obs_s[i];
    }
    if (obs_copy != nullptr && !cpy_vec)
      for (int i = tid; i < n; i += PP_THREADS) obs_copy[(size_t)env0 * O + i] = obs_s[i];
  }

  mark();
  // ---- the last CTA to finish turns the bitmap into the ascending id list (== reset_buf.nonzero(), fly.py:447)
  if (tid == 0) {
    s_last = (done_prev == gridDim.x - 1) ? 1 : 0;
    if (s_last) __threadfence();
  }
  __syncthreads();
  if (!s_last) return;
  // each thread owns a contiguous run of W masks: one sweep to count, one block scan, one sweep to write its ids
  const int nblk = (int)gridDim.x;
  const int W = (nblk + PP_THREADS - 1) / PP_THREADS;
  const int w_begin = min(tid * W, nblk), w_end = min(w_begin + W, nblk);
  int cnt = 0;
  for (int w = w_begin; w < w_end; ++w) cnt += __popcll(ld_status(bitmap + w));
  int incl = cnt;                                       // inclusive scan over the CTA
#pragma unroll
  for (int o2 = 1; o2 < 32; o2 <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o2);
    if (lane >= o2) incl += t;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  int warp_off = 0, base = 0;
#pragma unroll
  for (int i = 0; i < PP_WARPS; ++i) {
    if (i < warp) warp_off += s_scan[i];
    base += s_scan[i];
  }
  int pos = warp_off + incl - cnt;
  for (int w = w_begin; w < w_end; ++w) {
    unsigned long long word = ld_status(bitmap + w);
    while (word != 0ull) {
      const int b = __ffsll((long long)word) - 1;
      reset_ids[pos++] = w * ENVS + b;
      word &= word - 1ull;
    }
  }
  if (tid == 0) {
    *reset_count = base;
    scratch->done = 0u;                                 // re-armed for the next launch
  }
  if (tail.action_var != nullptr) {   // end-of-step housekeeping (ppo.py:236-237), once per launch
    const int steps = tail.steps > 0 ? tail.steps : 1;   // a fused rollout applies its T annealing steps at its end
    for (int i = tid; i < tail.num_act; i += PP_THREADS) {
      float v = tail.action_var[i];
      for (int j = 0; j < steps; ++j) v = fmaxf(tail.floor, v - tail.decay);   // iterated: bit-identical to T launches
      tail.action_var[i] = v;
    }
    if (tid == 0 && tail.step_counter != nullptr) *tail.step_counter += (unsigned long long)steps;
  }
}

// Pre-physics half of Fly.step (fly.py:626-660): one thread per (env, dof) element for the target vector,
// reset bookkeeping by the threads that own an env's first DOF.
__global__ void __launch_bounds__(256)
pre_physics_kernel(const __grid_constant__ FlybEnvCfg cfg, const float* __restrict__ actions,
                   float* __restrict__ actions_scaled, float* __restrict__ dof_targets,
                   long long* __restrict__ reset, long long* __restrict__ progress, float* __restrict__ potentials,
                   float* __restrict__ prev_potentials, float* __restrict__ root, float* __restrict__ dof_state) {
  __shared__ int s_slot[FLYB_MAX_DOF];  // dof -> action slot or -1
  const int D = cfg.num_dof, A = cfg.num_act;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    int slot = -1;
    for (int k = 0; k < A; ++k)
      if (cfg.action_idx[k] == d) slot = k;
    s_slot[d] = slot;
  }
  __syncthreads();
  const size_t total = (size_t)cfg.num_envs * D;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int env = (int)(i / D);
    const int d = (int)(i - (size_t)env * D);
    if (actions != nullptr) {
      const int k = s_slot[d];
      float tgt = cfg.initial_pos[d];                                  // fly.py:626
      if (k >= 0) {
        const float a = __ldg(actions + (size_t)env * A + k);
        tgt = scale_action(a, cfg.lower[d], cfg.upper[d]);             // fly.py:629-630
        actions_scaled[(size_t)env * A + k] = tgt;                     // fly.py:633-636
      }
      dof_targets[i] = tgt;                                            // fly.py:648-651
    }
    if (reset != nullptr && reset[env] != 0) {                       // fly.py:446-480
      reinterpret_cast<float2*>(dof_state)[i] = make_float2(cfg.initial_pos[d], cfg.initial_vel[d]);
      if (d < 13) root[(size_t)env * 13 + d] = cfg.origin_root[d];
      if (D < 13 && d == 0)
        for (int j = D; j < 13; ++j) root[(size_t)env * 13 + j] = cfg.origin_root[j];
    }
  }
  // the masks are cleared by pre_physics_finish_kernel: an env's DOF elements may span two CTAs, so the
  // bookkeeping has to wait for a launch boundary
}

__global__ void __launch_bounds__(256)
pre_physics_finish_kernel(const __grid_constant__ FlybEnvCfg cfg, long long* __restrict__ reset,
                          long long* __restrict__ progress, float* __restrict__ potentials,
                          float* __restrict__ prev_potentials) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= cfg.num_envs) return;
  if (reset[env] != 0) {
    const float tx = cfg.target[0] - cfg.origin_root[0];
    const float ty = cfg.target[1] - cfg.origin_root[1];
    const float p = -sqrtf(tx * tx + ty * ty + 0.0f) / cfg.dt;       // fly.py:470-473
    prev_potentials[env] = p;
    potentials[env] = p;
    reset[env] = 0;                                                  // fly.py:476-477
    progress[env] = 0;
  }
}

}  // namespace flyb

using namespace flyb;

extern "C" size_t flyb_env_scratch_bytes(int32_t num_envs) {
  // masks: one word per 64 envs (+ 1: the fused step publishes two words per 128-env tile); tile counts behind them
  return scan_scratch_bytes(num_envs > 0 ? num_envs : 1);
}

static int validate_cfg(const FlybEnvCfg* cfg) {
  FLYB_REQUIRE(cfg != nullptr, "cfg is NULL");
  FLYB_REQUIRE(cfg->num_envs > 0, "num_envs must be > 0 (got %d)", cfg->num_envs);
  FLYB_REQUIRE(cfg->num_dof > 0 && cfg->num_dof <= FLYB_MAX_DOF, "num_dof %d out of (0, %d]", cfg->num_dof, FLYB_MAX_DOF);
  FLYB_REQUIRE(cfg->num_act > 0 && cfg->num_act <= cfg->num_dof, "num_act %d out of (0, num_dof]", cfg->num_act);
  FLYB_REQUIRE(cfg->num_bodies > 0, "num_bodies must be > 0");
  for (int k = 0; k < cfg->num_act; ++k) {
    FLYB_REQUIRE(cfg->action_idx[k] >= 0 && cfg->action_idx[k] < cfg->num_dof, "action_idx[%d] out of range", k);
    FLYB_REQUIRE(k == 0 || cfg->action_idx[k] > cfg->action_idx[k - 1], "action_idx must be strictly ascending");
  }
  for (int b = 0; b < FLYB_NUM_ABDOMEN; ++b)
    FLYB_REQUIRE(cfg->abdomen_local[b] >= 0 && cfg->abdomen_local[b] < cfg->num_bodies, "abdomen_local[%d] out of range", b);
  for (int b = 0; b < FLYB_NUM_LEGTIPS; ++b)
    FLYB_REQUIRE(cfg->legtip_local[b] >= 0 && cfg->legtip_local[b] < cfg->num_bodies, "legtip_local[%d] out of range", b);
  return FLYB_OK;
}

extern "C" int flyb_env_post_physics(const FlybEnvCfg* cfg, const float* root, const float* dof_state,
                                     const float* force, const float* actions_scaled, float* potentials,
                                     float* prev_potentials, int64_t* progress, int64_t* reset, float* obs,
                                     float* reward, float* up_vec, float* heading_vec, int32_t* reset_ids,
                                     int32_t* reset_count, void* scratch, float* obs_copy, float* reward_copy,
                                     float* reward_sum, const FlybStepTail* tail, void* stream) {
  int rc = validate_cfg(cfg);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(root && dof_state && force && actions_scaled && potentials && prev_potentials && progress && reset &&
                   obs && reward && up_vec && heading_vec && reset_ids && reset_count && scratch,
               "flyb_env_post_physics: NULL buffer");
  FLYB_REQUIRE(aligned16(root) && aligned16(obs) && aligned16(dof_state),
               "flyb_env_post_physics: root/dof_state/obs must be 16-byte aligned");
  const int O = 19 + 3 * cfg->num_act;
  size_t smem = ((size_t)ENVS * (O + 13 + FORCE_VALS) * sizeof(float) + 15) / 16 * 16;
  // bulk-copied force blocks: measured SLOWER here than the gather (26.1 vs 21.9 us at 16384 envs, profiles/r02e): the
  // kernel's 256 small CTAs hide the gather's latency by occupancy, and 101 KB of shared memory per CTA takes that away.
  // In the fused step (one big CTA per SM, nothing to overlap with) the same copy wins.  Opt-in: FLYB_ENV_FORCE_BULK=1.
  static const bool want_bulk = [] { const char* ev = getenv("FLYB_ENV_FORCE_BULK"); return ev && ev[0] == '1'; }();
  const size_t fblk = (size_t)ENVS * cfg->num_bodies * 12;
  const int force_bulk = (want_bulk && aligned16(force) && fblk % 16 == 0 && smem + fblk <= 110 * 1024) ? 1 : 0;
  if (force_bulk) smem += fblk;
  rc = ensure_smem(reinterpret_cast<const void*>(post_physics_kernel), smem);
  if (rc != FLYB_OK) return rc;
  const int grid = (cfg->num_envs + ENVS - 1) / ENVS;
  FlybStepTail t{};
  if (tail && tail->action_var) {
    FLYB_REQUIRE(tail->num_act > 0 && tail->num_act <= FLYB_MAX_DOF, "flyb_env_post_physics: tail.num_act out of range");
    t = *tail;
  }
  post_physics_kernel<<<grid, PP_THREADS, smem, (cudaStream_t)stream>>>(
      *cfg, root, dof_state, force, actions_scaled, potentials, prev_potentials, (long long*)progress,
      (long long*)reset, obs, reward, up_vec, heading_vec, reset_ids, reset_count, (ScanScratch*)scratch, obs_copy,
      reward_copy, reward_sum, t, tc_get_trace(), force_bulk);
  FLYB_LAUNCH_CHECK("post_physics_kernel");
  return FLYB_OK;
}

// ---- flyb_env_post_physics for hosts that have no way to hand us scratch memory.
// Who that is: a binding that only forwards the simulator's own tensors (the ctypes stub of INTEGRATION.md run without
// torch or a CUDA-runtime binding has device pointers to Isaac Gym's wrapped views but no allocator of its own), and a
// host that calls us from inside the simulator's step loop, where a cudaMalloc would stall the whole device (it
// synchronises implicitly) or simply fail because the simulator reserved the card's memory up front.
// The reset-id compaction needs a counter and one mask word per 64 envs that are neither input nor output.  Here they
// live in the library image itself: every GPU that loads libflyb200.so carries these ~40 KB until the process exits.
constexpr size_t static_scratch_need(size_t n) {   // == scan_scratch_bytes(n), spelled as a constant expression
  return 16 + ((n + SCAN_ENVS - 1) / SCAN_ENVS + 1) * 8 + ((((n + 127) / 128 + 1) + 1) / 2) * 8;
}
constexpr size_t STATIC_SCRATCH_BYTES = (static_scratch_need(FLYB_STATIC_MAX_ENVS) + 255) / 256 * 256;
static __device__ __align__(256) unsigned char g_static_env_scratch[STATIC_SCRATCH_BYTES];

// In host code `g_static_env_scratch` is only a handle for the runtime's symbol table, NOT an address a kernel may
// dereference; the device address differs per GPU and is looked up once per device, with that device current.
static int static_env_scratch(void** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static void* dev_ptr[MAX_DEV] = {};
  int dev = -1;
  int rc = flyb::check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(dev >= 0 && dev < MAX_DEV, "flyb_env_post_physics_static: device ordinal %d out of [0, %d)", dev, MAX_DEV);
  std::lock_guard<std::mutex> lock(mu);
  if (dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    rc = flyb::check_cuda(cudaGetSymbolAddress(&p, g_static_env_scratch), "cudaGetSymbolAddress(static env scratch)");
    if (rc != FLYB_OK) return rc;
    FLYB_REQUIRE(p != nullptr, "flyb_env_post_physics_static: the runtime returned no address for the built-in scratch");
    dev_ptr[dev] = p;
  }
  *out = dev_ptr[dev];
  return FLYB_OK;
}

extern "C" int flyb_env_post_physics_static(const FlybEnvCfg* cfg, const float* root, const float* dof_state,
                                            const float* force, const float* actions_scaled, float* potentials,
                                            float* prev_potentials, int64_t* progress, int64_t* reset, float* obs,
                                            float* reward, float* up_vec, float* heading_vec, int32_t* reset_ids,
                                            int32_t* reset_count, float* obs_copy, float* reward_copy,
                                            float* reward_sum, const FlybStepTail* tail, void* stream) {
  int rc = validate_cfg(cfg);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg->num_envs <= FLYB_STATIC_MAX_ENVS,
               "flyb_env_post_physics_static: num_envs %d exceeds the built-in scratch (FLYB_STATIC_MAX_ENVS = %d); "
               "use flyb_env_post_physics with a caller-owned scratch buffer", cfg->num_envs, FLYB_STATIC_MAX_ENVS);
  FLYB_REQUIRE(scan_scratch_bytes(cfg->num_envs) <= STATIC_SCRATCH_BYTES,
               "flyb_env_post_physics_static: scratch layout grew past the built-in array");
  void* scratch = nullptr;
  rc = static_env_scratch(&scratch);   // first call on this device: symbol look-up (may load the module)
  if (rc != FLYB_OK) return rc;
  // The array still holds whatever the previous call (any env count, any stream) left.  The kernel relies on the
  // finished-CTA counter starting at zero; the mask words need no clearing (every CTA stores its own before counting).
  rc = flyb::check_cuda(cudaMemsetAsync(scratch, 0, 16, (cudaStream_t)stream), "cudaMemsetAsync(static env scratch header)");
  if (rc != FLYB_OK) return rc;
  return flyb_env_post_physics(cfg, root, dof_state, force, actions_scaled, potentials, prev_potentials, progress,
                               reset, obs, reward, up_vec, heading_vec, reset_ids, reset_count, scratch, obs_copy,
                               reward_copy, reward_sum, tail, stream);
}

extern "C" int flyb_env_pre_physics(const FlybEnvCfg* cfg, const float* actions, float* actions_scaled,
                                    float* dof_targets, int64_t* reset, int64_t* progress, float* potentials,
                                    float* prev_potentials, float* root, float* dof_state, void* stream) {
  int rc = validate_cfg(cfg);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(actions && actions_scaled && dof_targets, "flyb_env_pre_physics: NULL buffer");
  FLYB_REQUIRE(!reset || (progress && potentials && prev_potentials && root && dof_state),
               "flyb_env_pre_physics: the reset half needs progress, potentials, prev_potentials, root and dof_state");
  FLYB_REQUIRE(!reset || (reinterpret_cast<uintptr_t>(dof_state) & 7u) == 0, "dof_state must be 8-byte aligned");
  const size_t total = (size_t)cfg->num_envs * cfg->num_dof;
  const int grid = (int)((total + 255) / 256);
  pre_physics_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*cfg, actions, actions_scaled, dof_targets,
                                                            (long long*)reset, (long long*)progress, potentials,
                                                            prev_potentials, root, dof_state);
  FLYB_LAUNCH_CHECK("pre_physics_kernel");
  if (!reset) return FLYB_OK;   // low-gravity step order: Fly.reset() runs after simulate() (flyb_env_reset)
  pre_physics_finish_kernel<<<(cfg->num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      *cfg, (long long*)reset, (long long*)progress, potentials, prev_potentials);
  FLYB_LAUNCH_CHECK("pre_physics_finish_kernel");
  return FLYB_OK;
}

extern "C" int flyb_env_reset(const FlybEnvCfg* cfg, int64_t* reset, int64_t* progress, float* potentials,
                              float* prev_potentials, float* root, float* dof_state, void* stream) {
  int rc = validate_cfg(cfg);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(reset && progress && potentials && prev_potentials && root && dof_state, "flyb_env_reset: NULL buffer");
  FLYB_REQUIRE((reinterpret_cast<uintptr_t>(dof_state) & 7u) == 0, "dof_state must be 8-byte aligned");
  const size_t total = (size_t)cfg->num_envs * cfg->num_dof;
  pre_physics_kernel<<<(int)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      *cfg, nullptr, nullptr, nullptr, (long long*)reset, (long long*)progress, potentials, prev_potentials, root,
      dof_state);
  FLYB_LAUNCH_CHECK("pre_physics_kernel(reset)");
  pre_physics_finish_kernel<<<(cfg->num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      *cfg, (long long*)reset, (long long*)progress, potentials, prev_potentials);
  FLYB_LAUNCH_CHECK("pre_physics_finish_kernel");
  return FLYB_OK;
}
