// Kernel (b): policy act — Net.pi forward, Gaussian sample / log-prob / clip, and (fused) the pre-physics half
// of Fly.step: action scaling, DOF-target scatter and reset bookkeeping.
//
// Replaces reference ppo.py:210-220 (+ the all_obs rollout store, ppo.py:226) and fly.py:626-660, 446-480.
// The reference builds an 18x18 Cholesky factor and a MultivariateNormal every step (~112 aten launches);
// with a diagonal covariance that is exactly  a = mu + sqrt(var) * eps,
// logp = -0.5 * (A log 2pi + sum(((a - mu)/sqrt(var))^2)) - sum(log sqrt(var))   (torch.distributions' order).
#include <algorithm>

#include "env_math.cuh"
#include "mlp_simt.cuh"
#include "philox.cuh"
#include "tc_dispatch.cuh"

namespace flyb {

struct ActArgs {
  const float* obs;
  long long rows;
  const float* action_var;
  const float* noise;
  unsigned long long seed, offset;
  const unsigned long long* offset_dev;
  float* action;
  float* logp;
  float* mu_out;
  float* obs_store;
  // fused pre-physics (all NULL when has_env == 0)
  int has_env;
  float* actions_scaled;
  float* dof_targets;
  long long* reset;
  long long* progress;
  float* potentials;
  float* prev_potentials;
  float* root;
  float* dof_state;
};

template <int APAD>
__global__ void __launch_bounds__(MLP_THREADS, 1)
policy_act_kernel(const __grid_constant__ FlybNetLayout net, const float* __restrict__ params,
                  const __grid_constant__ ActArgs a, const __grid_constant__ FlybEnvCfg cfg) {
  extern __shared__ __align__(16) float smem[];
  const MlpSmem s = mlp_carve(smem, net.k0_pad);
  __shared__ float s_std[FLYB_MAX_DOF];
  __shared__ float s_half_log_det;
  __shared__ int s_slot[FLYB_MAX_DOF];
  const int A = net.num_act, O = net.num_obs;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x < A) s_std[threadIdx.x] = sqrtf(__ldg(a.action_var + threadIdx.x));
  if (a.has_env)
    for (int d = threadIdx.x; d < cfg.num_dof; d += MLP_THREADS) {
      int slot = -1;
      for (int k = 0; k < A; ++k)
        if (cfg.action_idx[k] == d) slot = k;
      s_slot[d] = slot;
    }
  __syncthreads();
  if (threadIdx.x == 0) {
    float hl = 0.0f;
    for (int k = 0; k < A; ++k) hl += logf(s_std[k]);
    s_half_log_det = hl;
  }

  const unsigned long long draw = a.offset + (a.offset_dev ? __ldg(a.offset_dev) : 0ull);
  const long long ntiles = (a.rows + MLP_TM - 1) / MLP_TM;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * MLP_TM;
    const int nrows = (int)min((long long)MLP_TM, a.rows - row0);
    mlp_load_obs_tile(s, a.obs, row0, nrows, O, net.k0_pad);
    if (a.obs_store) {  // rollout row all_obs[i] = obs (ppo.py:226)
      float* dst = a.obs_store + (size_t)row0 * O;
      for (int e = threadIdx.x; e < nrows * O; e += MLP_THREADS) {
        const int r = e / O, c = e - r * O;
        dst[e] = s.buf1[r * s.ld1 + c];
      }
    }
    mlp_forward_tile<true, false, APAD>(s, net, params);

    // ---- per-row epilogue: one warp per row
    const float a_log_2pi = (float)((double)A * 1.8378770664093453);   // k * log(2*pi)
    for (int r = warp; r < nrows; r += MLP_THREADS / 32) {
      const long long row = row0 + r;
      float* mu = s.buf0 + r * MLP_LD0 + MLP_COL_MU;   // overwritten in place by the scaled action
      float m = 0.0f;
      for (int k = lane; k < A; k += 32) {
        const float mu_k = mu[k];
        const float eps = a.noise ? __ldg(a.noise + (size_t)row * A + k)
                                  : philox_normal(a.seed, draw, (unsigned long long)row * A + k);
        const float sample = mu_k + s_std[k] * eps;                 // MultivariateNormal.sample (ppo.py:218)
        const float diff = (sample - mu_k) / s_std[k];
        m += diff * diff;
        const float clipped = fminf(fmaxf(sample, -1.0f), 1.0f);    // ppo.py:220 (after log_prob)
        a.action[(size_t)row * A + k] = clipped;
        if (a.mu_out) a.mu_out[(size_t)row * A + k] = mu_k;
        if (a.has_env) {
          const int d = cfg.action_idx[k];
          const float sc = scale_action(clipped, cfg.lower[d], cfg.upper[d]);   // fly.py:629-630
          a.actions_scaled[(size_t)row * A + k] = sc;                          // fly.py:633-636
          mu[k] = sc;
        }
      }
      m = warp_sum(m);
      if (lane == 0) a.logp[row] = -0.5f * (a_log_2pi + m) - s_half_log_det;  // ppo.py:219
      if (a.has_env) {
        __syncwarp();
        const int D = cfg.num_dof;
        const bool rst = a.reset != nullptr && a.reset[row] != 0;
        for (int d = lane; d < D; d += 32) {
          const int k = s_slot[d];
          a.dof_targets[(size_t)row * D + d] = (k >= 0) ? mu[k] : cfg.initial_pos[d];    // fly.py:626, 648-651
          if (rst)                                                                        // fly.py:454-455
            reinterpret_cast<float2*>(a.dof_state)[(size_t)row * D + d] =
                make_float2(cfg.initial_pos[d], cfg.initial_vel[d]);
        }
        if (rst) {                                                                        // fly.py:459-477
          if (lane < 13) a.root[(size_t)row * 13 + lane] = cfg.origin_root[lane];
          __syncwarp();
          if (lane == 0) {
            const float tx = cfg.target[0] - cfg.origin_root[0];
            const float ty = cfg.target[1] - cfg.origin_root[1];
            const float p = -sqrtf(tx * tx + ty * ty + 0.0f) / cfg.dt;
            a.prev_potentials[row] = p;
            a.potentials[row] = p;
            a.reset[row] = 0;
            a.progress[row] = 0;
          }
        }
      }
    }
    __syncthreads();
  }
}

struct FwdArgs {
  const float* obs;
  long long rows;
  float* mu_out;
  float* value_out;
};

template <int APAD, bool WANT_PI, bool WANT_V>
__global__ void __launch_bounds__(MLP_THREADS, 1)
net_forward_kernel(const __grid_constant__ FlybNetLayout net, const float* __restrict__ params,
                   const __grid_constant__ FwdArgs a) {
  extern __shared__ __align__(16) float smem[];
  const MlpSmem s = mlp_carve(smem, net.k0_pad);
  const int A = net.num_act, O = net.num_obs;
  const long long ntiles = (a.rows + MLP_TM - 1) / MLP_TM;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * MLP_TM;
    const int nrows = (int)min((long long)MLP_TM, a.rows - row0);
    mlp_load_obs_tile(s, a.obs, row0, nrows, O, net.k0_pad);
    mlp_forward_tile<WANT_PI, WANT_V, APAD>(s, net, params);
    if (WANT_PI)
      for (int e = threadIdx.x; e < nrows * A; e += MLP_THREADS) {
        const int r = e / A, k = e - r * A;
        a.mu_out[(size_t)row0 * A + e] = s.buf0[r * MLP_LD0 + MLP_COL_MU + k];
      }
    if (WANT_V)
      for (int r = threadIdx.x; r < nrows; r += MLP_THREADS) a.value_out[row0 + r] = s.buf0[r * MLP_LD0 + MLP_COL_VAL];
    __syncthreads();
  }
}

__global__ void anneal_var_kernel(float* var, int n, float decay, float floor_v, unsigned long long* counter) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) var[i] = fmaxf(floor_v, var[i] - decay);   // ppo.py:236-237
  if (i == 0 && counter) *counter += 1ull;
}

static int g_num_sms = 0;
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) g_num_sms = 148;
  }
  return g_num_sms;
}

template <typename K>
static int prepare(K kernel, size_t smem) {
  return ensure_smem(reinterpret_cast<const void*>(kernel), smem);
}

int validate_net(const FlybNetLayout* net, const float* params) {
  FLYB_REQUIRE(net != nullptr && params != nullptr, "net layout / params is NULL");
  FlybNetLayout ref;
  int rc = flyb_net_layout(net->num_obs, net->num_act, &ref);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(ref.total == net->total && ref.k0_pad == net->k0_pad, "FlybNetLayout was not produced by flyb_net_layout");
  FLYB_REQUIRE(aligned16(params), "params must be 16-byte aligned");
  return FLYB_OK;
}

// tensor-core path (policy_kernels_tc.cu)
}  // namespace flyb

using namespace flyb;

// Debug aid (not part of the reference-facing surface): device buffer of >= 160 u64 that CTA 0 of the tensor-core
// kernels fills with clock64() marks; NULL switches it off.
namespace flyb { void tc_set_trace(void* p); }
extern "C" int flyb_debug_set_trace(void* dev_buffer) {
  flyb::tc_set_trace(dev_buffer);
  return FLYB_OK;
}

extern "C" size_t flyb_wpack_bytes(const FlybNetLayout* net) { return net ? tc_wpack_bytes(net) : 0; }

extern "C" int flyb_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(wpack != nullptr && aligned16(wpack), "flyb_pack_weights: wpack must be a 16-byte aligned buffer");
  FLYB_REQUIRE(tc_supported(net), "flyb_pack_weights: the tensor-core path needs num_obs <= 160 and num_act <= 48");
  return tc_pack_weights(net, params, wpack, (cudaStream_t)stream);
}

extern "C" int flyb_env_step_fused_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg) {
  return (net && cfg && tc_step_supported(net, cfg)) ? 1 : 0;
}

extern "C" int flyb_env_step_fused(const FlybNetLayout* net, const float* params, const void* wpack,
                                   const FlybEnvCfg* cfg, const FlybStepIO* io, void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg && io && wpack, "flyb_env_step_fused: NULL cfg / io / wpack (the fused step runs on the tensor-core path only)");
  FLYB_REQUIRE(cfg->num_envs > 0 && cfg->num_dof > 0 && cfg->num_dof <= FLYB_MAX_DOF && cfg->num_bodies > 0,
               "flyb_env_step_fused: bad env cfg");
  FLYB_REQUIRE(cfg->num_act == net->num_act && 19 + 3 * cfg->num_act == net->num_obs,
               "flyb_env_step_fused: env cfg (A=%d) does not match the net (O=%d, A=%d)", cfg->num_act, net->num_obs,
               net->num_act);
  for (int k = 0; k < cfg->num_act; ++k)
    FLYB_REQUIRE(cfg->action_idx[k] >= 0 && cfg->action_idx[k] < cfg->num_dof, "action_idx[%d] out of range", k);
  for (int b = 0; b < FLYB_NUM_ABDOMEN; ++b)
    FLYB_REQUIRE(cfg->abdomen_local[b] >= 0 && cfg->abdomen_local[b] < cfg->num_bodies, "abdomen_local[%d] out of range", b);
  for (int b = 0; b < FLYB_NUM_LEGTIPS; ++b)
    FLYB_REQUIRE(cfg->legtip_local[b] >= 0 && cfg->legtip_local[b] < cfg->num_bodies, "legtip_local[%d] out of range", b);
  FLYB_REQUIRE(tc_step_supported(net, cfg), "flyb_env_step_fused: shape outside the fused path (needs num_obs <= 80, num_act <= 20)");
  FLYB_REQUIRE(io->root && io->dof_state && io->force && io->actions_scaled && io->potentials && io->prev_potentials &&
                   io->progress && io->reset && io->reward && io->up_vec && io->heading_vec && io->reset_ids &&
                   io->reset_count && io->scratch && io->action_var && io->action && io->logp && io->dof_targets,
               "flyb_env_step_fused: NULL buffer");
  FLYB_REQUIRE(aligned16(io->root) && aligned16(wpack) && (reinterpret_cast<uintptr_t>(io->dof_state) & 7u) == 0,
               "flyb_env_step_fused: root / wpack must be 16-byte, dof_state 8-byte aligned");
  FLYB_REQUIRE(io->anneal_steps >= 0, "flyb_env_step_fused: anneal_steps < 0");
  return tc_launch_step(net, params, wpack, cfg, io, (cudaStream_t)stream);
}

extern "C" int flyb_policy_act(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs,
                               int64_t rows,
                               const float* action_var, const float* noise, uint64_t seed, uint64_t offset,
                               const uint64_t* offset_dev, float* action, float* logp, float* mu_out, float* obs_store, const FlybEnvCfg* cfg,
                               float* actions_scaled, float* dof_targets, int64_t* reset, int64_t* progress,
                               float* potentials, float* prev_potentials, float* root, float* dof_state,
                               void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(obs && action_var && action && logp && rows > 0, "flyb_policy_act: NULL buffer or rows <= 0");
  ActArgs a{};
  a.obs = obs; a.rows = rows; a.action_var = action_var; a.noise = noise; a.seed = seed; a.offset = offset;
  a.offset_dev = (const unsigned long long*)offset_dev;
  a.action = action; a.logp = logp; a.mu_out = mu_out; a.obs_store = obs_store;
  FlybEnvCfg c{};
  if (cfg) {
    FLYB_REQUIRE(cfg->num_act == net->num_act && 19 + 3 * cfg->num_act == net->num_obs,
                 "flyb_policy_act: env cfg (A=%d) does not match the net (O=%d, A=%d)", cfg->num_act, net->num_obs,
                 net->num_act);
    FLYB_REQUIRE(cfg->num_envs == rows, "flyb_policy_act: rows (%lld) != cfg.num_envs (%d)", (long long)rows, cfg->num_envs);
    FLYB_REQUIRE(cfg->num_dof > 0 && cfg->num_dof <= FLYB_MAX_DOF, "num_dof out of range");
    FLYB_REQUIRE(actions_scaled && dof_targets, "flyb_policy_act: fused pre-physics needs actions_scaled and dof_targets");
    FLYB_REQUIRE(!reset || (progress && potentials && prev_potentials && root && dof_state),
                 "flyb_policy_act: the fused reset half needs progress, potentials, prev_potentials, root and dof_state");
    c = *cfg;
    a.has_env = 1;
    a.actions_scaled = actions_scaled; a.dof_targets = dof_targets; a.reset = (long long*)reset;
    a.progress = (long long*)progress; a.potentials = potentials; a.prev_potentials = prev_potentials;
    a.root = root; a.dof_state = dof_state;
  }
  if (wpack) {   // tensor-core path (3xTF32 on tcgen05, policy_kernels_tc.cu)
    FLYB_REQUIRE(tc_act_supported(net, cfg), "flyb_policy_act: wpack given but the tensor-core path does not cover "
                 "this shape (needs num_obs <= 160, num_act <= 48)");
    FLYB_REQUIRE(aligned16(wpack), "wpack must be 16-byte aligned");
    return tc_launch_act(net, params, wpack, obs, rows, action_var, noise, seed, offset, offset_dev, action, logp,
                         mu_out, obs_store, cfg, actions_scaled, dof_targets, reset, progress, potentials,
                         prev_potentials, root, dof_state, (cudaStream_t)stream);
  }
  const size_t smem = mlp_smem_bytes(net->k0_pad);
  const long long ntiles = (rows + MLP_TM - 1) / MLP_TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  if (net->num_act <= 32) {
    if ((rc = prepare(policy_act_kernel<32>, smem)) != FLYB_OK) return rc;
    policy_act_kernel<32><<<grid, MLP_THREADS, smem, (cudaStream_t)stream>>>(*net, params, a, c);
  } else {
    if ((rc = prepare(policy_act_kernel<64>, smem)) != FLYB_OK) return rc;
    policy_act_kernel<64><<<grid, MLP_THREADS, smem, (cudaStream_t)stream>>>(*net, params, a, c);
  }
  FLYB_LAUNCH_CHECK("policy_act_kernel");
  return FLYB_OK;
}

template <int APAD>
static int launch_forward(const FlybNetLayout* net, const float* params, const FwdArgs& a, int grid, size_t smem,
                          cudaStream_t st) {
  int rc;
  if (a.mu_out && a.value_out) {
    if ((rc = prepare(net_forward_kernel<APAD, true, true>, smem)) != FLYB_OK) return rc;
    net_forward_kernel<APAD, true, true><<<grid, MLP_THREADS, smem, st>>>(*net, params, a);
  } else if (a.mu_out) {
    if ((rc = prepare(net_forward_kernel<APAD, true, false>, smem)) != FLYB_OK) return rc;
    net_forward_kernel<APAD, true, false><<<grid, MLP_THREADS, smem, st>>>(*net, params, a);
  } else {
    if ((rc = prepare(net_forward_kernel<APAD, false, true>, smem)) != FLYB_OK) return rc;
    net_forward_kernel<APAD, false, true><<<grid, MLP_THREADS, smem, st>>>(*net, params, a);
  }
  return FLYB_OK;
}

extern "C" int flyb_net_forward(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs,
                                int64_t rows, float* mu_out, float* value_out, void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(obs && rows > 0 && (mu_out || value_out), "flyb_net_forward: NULL buffer or rows <= 0");
  if (wpack) {
    FLYB_REQUIRE(tc_supported(net) && aligned16(wpack), "flyb_net_forward: wpack given but the tensor-core path "
                 "does not cover this shape (needs num_obs <= 80, num_act <= 32) or wpack is misaligned");
    return tc_launch_forward(net, params, wpack, obs, rows, mu_out, value_out, (cudaStream_t)stream);
  }
  FwdArgs a{obs, rows, mu_out, value_out};
  const size_t smem = mlp_smem_bytes(net->k0_pad);
  const long long ntiles = (rows + MLP_TM - 1) / MLP_TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  rc = (net->num_act <= 32) ? launch_forward<32>(net, params, a, grid, smem, (cudaStream_t)stream)
                            : launch_forward<64>(net, params, a, grid, smem, (cudaStream_t)stream);
  if (rc != FLYB_OK) return rc;
  FLYB_LAUNCH_CHECK("net_forward_kernel");
  return FLYB_OK;
}

extern "C" int flyb_anneal_var(float* action_var, int32_t num_act, float decay, float floor_v,
                               uint64_t* step_counter, void* stream) {
  FLYB_REQUIRE(action_var && num_act > 0, "flyb_anneal_var: bad arguments");
  anneal_var_kernel<<<(num_act + 63) / 64, 64, 0, (cudaStream_t)stream>>>(action_var, num_act, decay, floor_v,
                                                                         (unsigned long long*)step_counter);
  FLYB_LAUNCH_CHECK("anneal_var_kernel");
  return FLYB_OK;
}
