// Kernel (b), tensor-core path: policy act (Net.pi forward + Gaussian sample / log-prob / clip + action scaling,
// DOF-target scatter and reset bookkeeping) and the plain Net.pi / Net.v forward, on tcgen05 / TMEM tiles fed by
// bulk TMA (mlp_tc.cuh).  Same results contract as the CUDA-core path in policy_kernels.cu (reference
// ppo.py:88-102, 210-220 and fly.py:626-660, 446-480); GEMMs are 3xTF32 with fp32 accumulation.
//
// Also here: flyb_pack_weights, which rewrites the flat fp32 parameters as the packed hi/lo weight stream the
// TMA producer reads (consumption order, 4 KB stages — layout in mlp_tc.cuh).
#include <algorithm>

#include "env_math.cuh"
#include "mlp_tc.cuh"
#include "philox.cuh"

namespace flyb {

int num_sms();
int validate_net(const FlybNetLayout* net, const float* params);

// ------------------------------------------------------------------------------------------------ weight packing
struct PackSeg {
  int first_stage, nstages;
  int N, ksteps;        // operand rows per slab, k-steps
  int w_off, ldw;       // source matrix inside the flat parameter buffer (row-major [rows][ldw])
  int n0, k0;           // first source row / column
  int n_rows, k_cols;   // valid source rows (from n0) / valid source columns (absolute)
  int transposed;       // 0: B[n][k] = W[n0 + n][k0 + k];  1: B[n][k] = W[k0 + k][n0 + n]  (backward operands)
};
constexpr int PACK_MAX_SEGS = 48;
struct PackPlan {
  PackSeg seg[PACK_MAX_SEGS];
  int nseg;
  int total_stages;
};

static PackPlan forward_plan(const FlybNetLayout& net) {
  using namespace tcm;
  PackPlan p{};
  int stage = 0;
  auto add = [&](int N, int ksteps, int w_off, int ldw, int n0, int k0, int n_rows, int k_cols) {
    PackSeg& s = p.seg[p.nseg++];
    s.first_stage = stage;
    s.nstages = ksteps * 2 * N * 32 / STAGE;
    s.N = N; s.ksteps = ksteps; s.w_off = w_off; s.ldw = ldw; s.n0 = n0; s.k0 = k0; s.n_rows = n_rows; s.k_cols = k_cols;
    s.transposed = 0;
    stage += s.nstages;
  };
  auto w1 = [&](int c) { add(CH, K0 / 8, net.off[0], net.k0_pad, c * CH, 0, CH, net.k0_pad); };
  auto w2 = [&](int c) { add(H2, CH / 8, net.off[2], H1, 0, c * CH, H2, H1); };
  // issue order of the MMA warp: W1[0], W1[1], then (W2[c], W1[c+2]) ...
  w1(0);
  w1(1);
  for (int c = 0; c < NCH; ++c) {
    w2(c);
    if (c + 2 < NCH) w1(c + 2);
  }
  add(H3, H2 / 8, net.off[4], H2, 0, 0, H3, H2);                   // W3pi  (to_mean.0)
  add(H3, H2 / 8, net.off[8], H2, 0, 0, H3, H2);                   // W3v   (to_value.0)
  add(APAD, H3 / 8, net.off[6], H3, 0, 0, net.num_act, H3);        // W4    (to_mean.2), rows >= A are zero
  p.total_stages = stage;
  return p;
}

__global__ void __launch_bounds__(256)
pack_weights_kernel(const __grid_constant__ PackPlan plan, const float* __restrict__ params,
                    float* __restrict__ wpack) {
  const int total = plan.total_stages * (tcm::STAGE / 4);
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int stage = idx / (tcm::STAGE / 4);
    int si = 0;
    for (int i = 1; i < plan.nseg; ++i)
      if (stage >= plan.seg[i].first_stage) si = i;
    const PackSeg& s = plan.seg[si];
    const int li = idx - s.first_stage * (tcm::STAGE / 4);
    const int slab_floats = s.N * 8;
    const int slab = li / slab_floats, e = li - slab * slab_floats;
    const int ks = slab >> 1, is_lo = slab & 1;
    const int kc = e / (s.N * 4), rem = e - kc * (s.N * 4);
    const int n = rem >> 2, kk = rem & 3;
    const int k = s.k0 + ks * 8 + kc * 4 + kk;
    float w = 0.0f;
    if (n < s.n_rows && k < s.k_cols)
      w = s.transposed ? params[s.w_off + (size_t)k * s.ldw + s.n0 + n] : params[s.w_off + (size_t)(s.n0 + n) * s.ldw + k];
    const float hi = __uint_as_float(tc::tf32_rna_bits(w));
    wpack[idx] = is_lo ? __uint_as_float(tc::tf32_rna_bits(w - hi)) : hi;
  }
}

// ------------------------------------------------------------------------------------------------ finalisers
// Cooperative copy of a staged tile to global memory: one bulk TMA store when size and address allow, plain
// coalesced stores otherwise.  Called by all 256 epilogue threads after the staging writes were fenced + synced.
FLYB_D void store_tile(float* dst, const float* src_smem, int nfloats) {
  const uint32_t bytes = (uint32_t)nfloats * 4u;
  if ((bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
    if (threadIdx.x == 0) tc::tma_store_1d(dst, src_smem, bytes);
  } else {
    for (int i = threadIdx.x; i < nfloats; i += tcm::EPI_THREADS) dst[i] = src_smem[i];
  }
}

struct FwdFinal {
  float* mu_out;
  float* value_out;
  int A;
  FLYB_D void begin_tile(const tcm::Smem&, long long, int, bool) {}
  FLYB_D void obs_consumed(const tcm::Smem& s, bool) {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tc::mbar_arrive(s.obs_empty);
  }
  FLYB_D void value(long long row0, int r, int nrows, float v) {
    if (r < nrows) value_out[row0 + r] = v;
  }
  FLYB_D void pi(const tcm::Smem& s, long long row0, int r, int h, int nrows, const float (&mu)[tcm::APAD]) {
    if (threadIdx.x == 0) tc::tma_store_wait_read<0>();
    tcm::epi_bar_sync();
    if (h == 0) {
#pragma unroll
      for (int k = 0; k < tcm::APAD; ++k)
        if (k < A) s.out[r * A + k] = mu[k];
    }
    tc::fence_proxy_async();
    tcm::epi_bar_sync();
    store_tile(mu_out + (size_t)row0 * A, s.out, nrows * A);
    if (threadIdx.x == 0) tc::tma_store_commit();
  }
  FLYB_D void finish(const tcm::Smem&) {
    if (threadIdx.x == 0) tc::tma_store_wait<0>();
  }
};

struct ActFinal {
  // act
  const float* action_var;
  const float* noise;
  unsigned long long seed, draw;
  float* action;
  float* logp;
  float* mu_out;
  float* obs_store;
  int A, O;
  // fused pre-physics
  const FlybEnvCfg* cfg;   // points at the kernel's __grid_constant__ copy
  int has_env;
  float* actions_scaled;
  float* dof_targets;
  long long* reset;
  long long* progress;
  float* potentials;
  float* prev_potentials;
  float* root;
  float* dof_state;
  // shared scratch owned by the kernel
  const float* s_std;      // [A]
  const int* s_slot;       // [D] dof -> action slot or -1
  float half_log_det;

  FLYB_D void begin_tile(const tcm::Smem& s, long long row0, int nrows, bool) {
    if (obs_store) {   // rollout row all_obs[i] = obs (ppo.py:226)
      tc::fence_proxy_async();
      store_tile(obs_store + (size_t)row0 * O, s.obs, nrows * O);
      if (threadIdx.x == 0) tc::tma_store_commit();
    }
  }
  FLYB_D void obs_consumed(const tcm::Smem& s, bool) {
    if (obs_store && threadIdx.x == 0) tc::tma_store_wait_read<0>();   // the bulk store has read the staging buffer
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tc::mbar_arrive(s.obs_empty);
  }
  FLYB_D void value(long long, int, int, float) {}

  FLYB_D void pi(const tcm::Smem& s, long long row0, int r, int h, int nrows, const float (&mu)[tcm::APAD]) {
    const int D = has_env ? cfg->num_dof : 0;
    float* act_s = s.out;                 // [TM][A]  mu, then the clipped action
    float* scaled_s = act_s + tcm::TM * A;  // [TM][A]
    float* tgt_s = scaled_s + tcm::TM * A;  // [TM][D]
    const long long row = row0 + r;
    const bool valid = r < nrows;
    if (threadIdx.x == 0) tc::tma_store_wait_read<0>();   // previous tile's bulk stores have drained the staging
    tcm::epi_bar_sync();
    if (h == 0) {
#pragma unroll
      for (int k = 0; k < tcm::APAD; ++k)
        if (k < A) act_s[r * A + k] = mu[k];
    }
    tcm::epi_bar_sync();
    // both column halves sample their share of the A actions of the row
    const int half = (A + 1) >> 1;
    const int k_lo = h ? half : 0, k_hi = h ? A : half;
    float m = 0.0f;
    if (valid) {
      for (int k = k_lo; k < k_hi; ++k) {
        const float mu_k = act_s[r * A + k];
        const float eps = noise ? __ldg(noise + (size_t)row * A + k)
                                : philox_normal(seed, draw, (unsigned long long)row * A + k);
        const float sample = mu_k + s_std[k] * eps;                 // MultivariateNormal.sample (ppo.py:218)
        const float diff = (sample - mu_k) / s_std[k];
        m += diff * diff;
        const float clipped = fminf(fmaxf(sample, -1.0f), 1.0f);    // ppo.py:220 (after log_prob)
        act_s[r * A + k] = clipped;
        if (mu_out) mu_out[(size_t)row * A + k] = mu_k;
        if (has_env) {
          const int d = cfg->action_idx[k];
          scaled_s[r * A + k] = scale_action(clipped, cfg->lower[d], cfg->upper[d]);   // fly.py:629-630
        }
      }
    }
    if (h == 1) s.vpart[r] = m;
    tcm::epi_bar_sync();
    if (h == 0 && valid) {
      const float a_log_2pi = (float)((double)A * 1.8378770664093453);
      logp[row] = -0.5f * (a_log_2pi + (m + s.vpart[r])) - half_log_det;               // ppo.py:219
    }
    if (has_env && valid) {
      const bool rst = reset[row] != 0;
      const int dh = (D + 1) >> 1;
      const int d_lo = h ? dh : 0, d_hi = h ? D : dh;
      for (int d = d_lo; d < d_hi; ++d) {
        const int k = s_slot[d];
        tgt_s[r * D + d] = (k >= 0) ? scaled_s[r * A + k] : cfg->initial_pos[d];       // fly.py:626, 648-651
        if (rst)                                                                        // fly.py:454-455
          reinterpret_cast<float2*>(dof_state)[(size_t)row * D + d] = make_float2(cfg->initial_pos[d], cfg->initial_vel[d]);
      }
      if (rst) {                                                                        // fly.py:459-477
        if (h == 1) {
#pragma unroll
          for (int j = 0; j < 13; ++j) root[(size_t)row * 13 + j] = cfg->origin_root[j];
        } else {
          const float tx = cfg->target[0] - cfg->origin_root[0];
          const float ty = cfg->target[1] - cfg->origin_root[1];
          const float p = -sqrtf(tx * tx + ty * ty + 0.0f) / cfg->dt;
          prev_potentials[row] = p;
          potentials[row] = p;
          progress[row] = 0;
        }
      }
    }
    tc::fence_proxy_async();
    tcm::epi_bar_sync();
    // the reset mask is cleared only after both halves of the row have read it
    if (has_env && valid && h == 0 && reset[row] != 0) reset[row] = 0;
    store_tile(action + (size_t)row0 * A, act_s, nrows * A);
    if (has_env) {
      store_tile(actions_scaled + (size_t)row0 * A, scaled_s, nrows * A);              // fly.py:633-636
      store_tile(dof_targets + (size_t)row0 * D, tgt_s, nrows * D);
    }
    if (threadIdx.x == 0) tc::tma_store_commit();
  }
  FLYB_D void finish(const tcm::Smem&) {
    if (threadIdx.x == 0) tc::tma_store_wait<0>();
  }
};

// ------------------------------------------------------------------------------------------------ kernels
template <bool WANT_PI, bool WANT_V>
__global__ void __launch_bounds__(tcm::THREADS, 1)
net_forward_tc_kernel(const __grid_constant__ FlybNetLayout net, const __grid_constant__ tcm::FwdArgs a,
                      float* __restrict__ mu_out, float* __restrict__ value_out) {
  extern __shared__ unsigned char smem_raw[];
  FwdFinal fin{mu_out, value_out, net.num_act};
  tcm::forward_persistent<WANT_PI, WANT_V>(net, a, fin, smem_raw);
}

struct ActTcArgs {
  const float* action_var;
  const float* noise;
  unsigned long long seed, offset;
  const unsigned long long* offset_dev;
  float* action;
  float* logp;
  float* mu_out;
  float* obs_store;
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

__global__ void __launch_bounds__(tcm::THREADS, 1)
policy_act_tc_kernel(const __grid_constant__ FlybNetLayout net, const __grid_constant__ tcm::FwdArgs a,
                     const __grid_constant__ ActTcArgs x, const __grid_constant__ FlybEnvCfg cfg) {
  extern __shared__ unsigned char smem_raw[];
  __shared__ float s_std[FLYB_MAX_DOF];
  __shared__ int s_slot[FLYB_MAX_DOF];
  __shared__ float s_half_log_det;
  const int A = net.num_act;
  if (threadIdx.x < A) s_std[threadIdx.x] = sqrtf(__ldg(x.action_var + threadIdx.x));
  if (x.has_env)
    for (int d = threadIdx.x; d < cfg.num_dof; d += tcm::THREADS) {
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
  __syncthreads();
  ActFinal fin;
  fin.action_var = x.action_var; fin.noise = x.noise; fin.seed = x.seed;
  fin.draw = x.offset + (x.offset_dev ? __ldg(x.offset_dev) : 0ull);
  fin.action = x.action; fin.logp = x.logp; fin.mu_out = x.mu_out; fin.obs_store = x.obs_store;
  fin.A = A; fin.O = net.num_obs;
  fin.cfg = &cfg; fin.has_env = x.has_env;
  fin.actions_scaled = x.actions_scaled; fin.dof_targets = x.dof_targets; fin.reset = x.reset; fin.progress = x.progress;
  fin.potentials = x.potentials; fin.prev_potentials = x.prev_potentials; fin.root = x.root; fin.dof_state = x.dof_state;
  fin.s_std = s_std; fin.s_slot = s_slot; fin.half_log_det = s_half_log_det;
  tcm::forward_persistent<true, false>(net, a, fin, smem_raw);
}

// ------------------------------------------------------------------------------------------------ host side
bool tc_supported(const FlybNetLayout* net) { return net->k0_pad == tcm::K0 && net->num_act <= tcm::APAD; }

size_t tc_wpack_bytes(const FlybNetLayout* net) {
  if (!tc_supported(net)) return 0;
  return (size_t)forward_plan(*net).total_stages * tcm::STAGE;
}

int tc_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, cudaStream_t st) {
  const PackPlan plan = forward_plan(*net);
  const int total = plan.total_stages * (tcm::STAGE / 4);
  pack_weights_kernel<<<std::min((total + 255) / 256, 2 * 148), 256, 0, st>>>(plan, params, (float*)wpack);
  FLYB_LAUNCH_CHECK("pack_weights_kernel");
  return FLYB_OK;
}

template <typename K>
static int prepare(K kernel, size_t smem) {
  return ensure_smem(reinterpret_cast<const void*>(kernel), smem);
}

int tc_launch_forward(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs,
                      long long rows, float* mu_out, float* value_out, cudaStream_t st) {
  tcm::FwdArgs a{obs, (const unsigned char*)wpack, params, rows, net->num_obs, net->num_act,
                 mu_out ? (size_t)tcm::TM * net->num_act * 4 : 0};
  const size_t smem = tcm::smem_bytes(a.out_bytes);
  const long long ntiles = (rows + tcm::TM - 1) / tcm::TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  int rc;
  if (mu_out && value_out) {
    if ((rc = prepare(net_forward_tc_kernel<true, true>, smem)) != FLYB_OK) return rc;
    net_forward_tc_kernel<true, true><<<grid, tcm::THREADS, smem, st>>>(*net, a, mu_out, value_out);
  } else if (mu_out) {
    if ((rc = prepare(net_forward_tc_kernel<true, false>, smem)) != FLYB_OK) return rc;
    net_forward_tc_kernel<true, false><<<grid, tcm::THREADS, smem, st>>>(*net, a, mu_out, value_out);
  } else {
    if ((rc = prepare(net_forward_tc_kernel<false, true>, smem)) != FLYB_OK) return rc;
    net_forward_tc_kernel<false, true><<<grid, tcm::THREADS, smem, st>>>(*net, a, mu_out, value_out);
  }
  FLYB_LAUNCH_CHECK("net_forward_tc_kernel");
  return FLYB_OK;
}

// true when the fused act + pre-physics staging fits next to the weight ring
bool tc_act_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg) {
  if (!tc_supported(net)) return false;
  const size_t out = (size_t)tcm::TM * (2 * net->num_act + (cfg ? cfg->num_dof : 0)) * 4;
  return tcm::smem_bytes(out) <= 227u * 1024u;
}

int tc_launch_act(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs, long long rows,
                  const float* action_var, const float* noise, uint64_t seed, uint64_t offset,
                  const uint64_t* offset_dev, float* action, float* logp, float* mu_out, float* obs_store,
                  const FlybEnvCfg* cfg, float* actions_scaled, float* dof_targets, int64_t* reset, int64_t* progress,
                  float* potentials, float* prev_potentials, float* root, float* dof_state, cudaStream_t st) {
  const size_t out = (size_t)tcm::TM * (2 * net->num_act + (cfg ? cfg->num_dof : 0)) * 4;
  tcm::FwdArgs a{obs, (const unsigned char*)wpack, params, rows, net->num_obs, net->num_act, out};
  ActTcArgs x{};
  x.action_var = action_var; x.noise = noise; x.seed = seed; x.offset = offset;
  x.offset_dev = (const unsigned long long*)offset_dev;
  x.action = action; x.logp = logp; x.mu_out = mu_out; x.obs_store = obs_store;
  FlybEnvCfg c{};
  if (cfg) {
    c = *cfg;
    x.has_env = 1;
    x.actions_scaled = actions_scaled; x.dof_targets = dof_targets; x.reset = (long long*)reset;
    x.progress = (long long*)progress; x.potentials = potentials; x.prev_potentials = prev_potentials;
    x.root = root; x.dof_state = dof_state;
  }
  const size_t smem = tcm::smem_bytes(out);
  const long long ntiles = (rows + tcm::TM - 1) / tcm::TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  int rc;
  if ((rc = prepare(policy_act_tc_kernel, smem)) != FLYB_OK) return rc;
  policy_act_tc_kernel<<<grid, tcm::THREADS, smem, st>>>(*net, a, x, c);
  FLYB_LAUNCH_CHECK("policy_act_tc_kernel");
  return FLYB_OK;
}

}  // namespace flyb
