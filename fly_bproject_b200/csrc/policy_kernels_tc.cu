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
  // ---- backward operands (ppo_kernels_tc.cu): B[n][k] = W[k][n], consumed after the forward stages
  auto addt = [&](int N, int ksteps, int w_off, int ldw, int n0, int n_rows, int k_cols) {
    add(N, ksteps, w_off, ldw, n0, 0, n_rows, k_cols);
    p.seg[p.nseg - 1].transposed = 1;
  };
  addt(H3, APAD / 8, net.off[6], H3, 0, H3, net.num_act);          // W4^T : dM1 = dZmu * W4       (k = action)
  addt(H2, H3 / 8, net.off[4], H2, 0, H2, H3);                     // W3pi^T: dH2 += dZm1 * W3pi    (k = m1 feature)
  addt(H2, H3 / 8, net.off[8], H2, 0, H2, H3);                     // W3v^T : dH2 += dZv1 * W3v
  addt(128, H2 / 8, net.off[2], H1, 0, 128, H2);                   // W2^T, h1 features [0,128)   (k = h2 feature)
  addt(128, H2 / 8, net.off[2], H1, 128, 128, H2);                 // W2^T, h1 features [128,256)
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

// Inverse of the mapping above: for every parameter, where its hi copies sit in the weight stream (at most two: the
// forward operand and the transposed backward operand) and how far behind the lo copy follows.  Built once by
// flyb_pack_weights; ppo_adam_kernel uses it to refresh the stream in place while it updates the parameters.
// Entry i = {dst0, delta0, dst1, delta1} in floats, dst = -1: none.
__global__ void __launch_bounds__(256)
pack_index_kernel(const __grid_constant__ PackPlan plan, int* __restrict__ inv) {
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
    if (slab & 1) continue;   // lo slab: reached through delta
    const int ks = slab >> 1;
    const int kc = e / (s.N * 4), rem = e - kc * (s.N * 4);
    const int n = rem >> 2, kk = rem & 3;
    const int k = s.k0 + ks * 8 + kc * 4 + kk;
    if (!(n < s.n_rows && k < s.k_cols)) continue;
    const int src = s.transposed ? s.w_off + k * s.ldw + s.n0 + n : s.w_off + (s.n0 + n) * s.ldw + k;
    int* ent = inv + 4 * (size_t)src;
    if (atomicCAS(ent + 0, -1, idx) == -1) ent[1] = slab_floats;
    else if (atomicCAS(ent + 2, -1, idx) == -1) ent[3] = slab_floats;
    else __trap();            // a third copy would need a wider table: the plan above never makes one
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
  FLYB_D void setup(int) {}
  FLYB_D void begin_tile(const tcm::Smem&, long long, int, int) {}
  FLYB_D void obs_consumed(const tcm::Smem& s) {
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tc::mbar_arrive(s.obs_empty);
  }
  FLYB_D void value(long long row0, int r, int nrows, float v) {
    if (r < nrows) value_out[row0 + r] = v;
  }
  FLYB_D void early(const tcm::Smem&, long long, int) {}
  FLYB_D void pre_pi(const tcm::Smem&, long long, int, int, int) {
    if (threadIdx.x == 0) tc::tma_store_wait_read<0>();   // previous tile's bulk store has drained the staging
    tcm::epi_bar_sync();
  }
  FLYB_D void pi(const tcm::Smem& s, long long row0, int r, int h, int nrows, const float (&mu)[tcm::APAD]) {
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

// per-kernel tables the act finaliser reads from shared memory
struct ActTables {
  float std[FLYB_MAX_DOF];      // sqrt(action_var[k])
  float inv_std[FLYB_MAX_DOF];
  int dof[FLYB_MAX_DOF];        // action k -> DOF it drives (cfg.action_idx)
  float log_std[FLYB_MAX_DOF];
  float lo[FLYB_MAX_DOF];       // lower limit of the DOF action k drives
  float hi[FLYB_MAX_DOF];
  float init_pos[FLYB_MAX_DOF]; // per DOF
  float init_vel[FLYB_MAX_DOF];
  int slot[FLYB_MAX_DOF];       // dof -> action slot or -1
  float half_log_det;
};

// Act finaliser.  The 256 epilogue threads work on the tile's (rows x A) actions as ONE flat array, a pair of
// consecutive elements per thread and trip (balanced, conflict-free shared memory, coalesced float2 global stores);
// per-row quantities (log-prob, reset bookkeeping) are finished by the thread that owns the row.
// Staging in s.out (floats): mu_s [TM*A] | eps_s, later scaled_s [TM*A] | dsq_s [TM*A].
struct ActFinal {
  const float* noise;
  unsigned long long seed, draw;
  float* action;
  float* logp;
  float* mu_out;
  float* obs_store;
  int A, O, D;
  // fused pre-physics
  const FlybEnvCfg* cfg;   // the kernel's __grid_constant__ copy
  int has_env;
  float* actions_scaled;
  float* dof_targets;
  long long* reset;
  long long* progress;
  float* potentials;
  float* prev_potentials;
  float* root;
  float* dof_state;
  const ActTables* t;      // shared memory
  ActTables* tw;           // the same table, writable by setup()
  const float* action_var;
  long long rst_raw;       // this row's reset flag, fetched at the start of the tile (consumed at its end)
  unsigned long long* trace;   // debug timeline (slots 128..159), thread 0 of CTA 0
  int ti;
  FLYB_D void mark() {
    if (trace != nullptr && threadIdx.x == 0 && blockIdx.x == 0 && ti < 160) trace[ti++] = clock64();
  }

  // per-kernel tables, built by the 32 lanes of the producer warp while the first tile is being staged
  FLYB_D void setup(int lane) {
    ActTables& tab = *tw;
    for (int k = lane; k < A; k += 32) {
      tab.std[k] = sqrtf(__ldg(action_var + k));
      tab.inv_std[k] = 1.0f / tab.std[k];
      tab.log_std[k] = logf(tab.std[k]);
      if (has_env) {
        const int d = cfg->action_idx[k];
        tab.dof[k] = d;
        tab.lo[k] = cfg->lower[d];
        tab.hi[k] = cfg->upper[d];
      }
    }
    if (has_env)
      for (int d = lane; d < cfg->num_dof; d += 32) {
        int slot = -1;
        for (int k = 0; k < A; ++k)
          if (cfg->action_idx[k] == d) slot = k;
        tab.slot[d] = slot;
        tab.init_pos[d] = cfg->initial_pos[d];
        tab.init_vel[d] = cfg->initial_vel[d];
      }
    __syncwarp();
    if (lane == 0) {
      float hl = 0.0f;
      for (int k = 0; k < A; ++k) hl += tab.log_std[k];   // same summation order as the CUDA-core path
      tab.half_log_det = hl;
    }
    __syncwarp();
  }
  FLYB_D void begin_tile(const tcm::Smem& s, long long row0, int r, int nrows) {
    rst_raw = (has_env && reset != nullptr && r < nrows) ? reset[row0 + r] : 0;
    if (obs_store) {   // rollout row all_obs[i] = obs (ppo.py:226)
      store_tile(obs_store + (size_t)row0 * O, s.obs, nrows * O);
      if (threadIdx.x == 0) tc::tma_store_commit();
    }
  }
  FLYB_D void obs_consumed(const tcm::Smem& s) {
    if (obs_store && threadIdx.x == 0) tc::tma_store_wait_read<0>();   // the bulk store has read the staging buffer
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tc::mbar_arrive(s.obs_empty);
  }
  FLYB_D void value(long long, int, int, float) {}

  // eps for the whole tile -> eps_s; independent of mu, so it runs while the first MMAs of the tile do
  FLYB_D void early(const tcm::Smem& s, long long row0, int nrows) {
    float* eps_s = s.out + tcm::TM * A;
    tcm::epi_bar_sync();   // every thread has left the previous tile's pi(), which reads this staging
    const int n = nrows * A;
    const unsigned long long g0 = (unsigned long long)row0 * A;   // even: row0 is a multiple of 128
    for (int e = 2 * threadIdx.x; e < n; e += 2 * tcm::EPI_THREADS) {
      float n0, n1;
      if (noise) {
        n0 = __ldg(noise + g0 + e);
        n1 = (e + 1 < n) ? __ldg(noise + g0 + e + 1) : 0.0f;
      } else {
        philox_normal_pair(seed, draw, (g0 + e) >> 1, n0, n1);
      }
      eps_s[e] = n0;
      if (e + 1 < n) eps_s[e + 1] = n1;
    }
  }
  FLYB_D void pre_pi(const tcm::Smem&, long long, int, int, int) {}
  FLYB_D void pi(const tcm::Smem& s, long long row0, int r, int h, int nrows, const float (&mu)[tcm::APAD]) {
    float* mu_s = s.out;
    float* sc_s = mu_s + tcm::TM * A;    // eps (written by early()), overwritten by the scaled action
    float* dsq_s = sc_s + tcm::TM * A;   // ((a - mu) / std)^2 per element
    if (h == 0) {
#pragma unroll
      for (int k = 0; k < tcm::APAD; ++k)
        if (k < A) mu_s[r * A + k] = mu[k];
    }
    mark();
    tcm::epi_bar_sync();
    mark();
    const int n = nrows * A;
    const size_t g0 = (size_t)row0 * A;
    const int kstep = (2 * tcm::EPI_THREADS) % A;
    // float2 stores need 8-byte aligned rows: all_acts[i] starts at i * N * A floats (odd for odd N and odd A)
    const bool pair_ok = ((reinterpret_cast<uintptr_t>(action) | (has_env ? reinterpret_cast<uintptr_t>(actions_scaled) : 0)) & 7u) == 0;
    int e = 2 * threadIdx.x;
    int k = e % A;
    for (; e < n; e += 2 * tcm::EPI_THREADS) {
      float cl[2], sc[2] = {0.0f, 0.0f};
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int ee = (e + u < n) ? e + u : e;
        int kk = k + u;
        if (kk >= A) kk -= A;
        const float mu_k = mu_s[ee];
        const float sample = mu_k + t->std[kk] * sc_s[ee];            // MultivariateNormal.sample (ppo.py:218)
        const float diff = (sample - mu_k) * t->inv_std[kk];
        cl[u] = fminf(fmaxf(sample, -1.0f), 1.0f);                    // ppo.py:220 (after log_prob)
        if (has_env) sc[u] = scale_action(cl[u], t->lo[kk], t->hi[kk]);   // fly.py:629-630
        if (e + u < n) {
          dsq_s[ee] = diff * diff;
          if (has_env) sc_s[ee] = sc[u];
          if (mu_out) mu_out[g0 + ee] = mu_k;
        }
      }
      if (e + 1 < n && !pair_ok) {
        action[g0 + e] = cl[0];
        action[g0 + e + 1] = cl[1];
        if (has_env) { actions_scaled[g0 + e] = sc[0]; actions_scaled[g0 + e + 1] = sc[1]; }
      } else if (e + 1 < n) {
        *reinterpret_cast<float2*>(action + g0 + e) = make_float2(cl[0], cl[1]);
        if (has_env) *reinterpret_cast<float2*>(actions_scaled + g0 + e) = make_float2(sc[0], sc[1]);   // fly.py:633-636
      } else {
        action[g0 + e] = cl[0];
        if (has_env) actions_scaled[g0 + e] = sc[0];
      }
      k += kstep;
      if (k >= A) k -= A;
    }
    mark();
    tcm::epi_bar_sync();
    mark();
    const long long row = row0 + r;
    if (h == 0 && r < nrows) {
      float m = 0.0f;
      for (int kk = 0; kk < A; ++kk) m += dsq_s[r * A + kk];
      const float a_log_2pi = (float)((double)A * 1.8378770664093453);
      logp[row] = -0.5f * (a_log_2pi + m) - t->half_log_det;          // ppo.py:219
    }
    mark();
    if (has_env) {
      // DOF position targets, flat over (rows x D): the driven DOFs take the scaled action, the others initial_pos
      const int nd = nrows * D;
      const int rstep = tcm::EPI_THREADS / D, dstep = tcm::EPI_THREADS % D;
      int rr = threadIdx.x / D, d = threadIdx.x - rr * D;
      float* tg = dof_targets + (size_t)row0 * D;
      for (int f = threadIdx.x; f < nd; f += tcm::EPI_THREADS) {
        const int slot = t->slot[d];
        tg[f] = (slot >= 0) ? sc_s[rr * A + slot] : t->init_pos[d];   // fly.py:626, 648-651
        rr += rstep;
        d += dstep;
        if (d >= D) { d -= D; ++rr; }
      }
      if (h == 1 && r < nrows && rst_raw != 0) {                      // Fly.reset (fly.py:446-480), rare
        for (int dd = 0; dd < D; ++dd)
          reinterpret_cast<float2*>(dof_state)[(size_t)row * D + dd] = make_float2(t->init_pos[dd], t->init_vel[dd]);
#pragma unroll
        for (int j = 0; j < 13; ++j) root[(size_t)row * 13 + j] = cfg->origin_root[j];
        const float tx = cfg->target[0] - cfg->origin_root[0];
        const float ty = cfg->target[1] - cfg->origin_root[1];
        const float p = -sqrtf(tx * tx + ty * ty + 0.0f) / cfg->dt;
        prev_potentials[row] = p;
        potentials[row] = p;
        progress[row] = 0;
        reset[row] = 0;
      }
    }
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
  FwdFinal fin{mu_out, value_out, net.num_act};
  tcm::forward_persistent<WANT_PI, WANT_V>(net, a, fin);
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
  __shared__ ActTables tab;
  const int A = net.num_act;
  if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[190] = clock64();   // kernel entry
  ActFinal fin;
  fin.noise = x.noise; fin.seed = x.seed;
  fin.draw = x.offset + (x.offset_dev ? __ldg(x.offset_dev) : 0ull);
  fin.action = x.action; fin.logp = x.logp; fin.mu_out = x.mu_out; fin.obs_store = x.obs_store;
  fin.A = A; fin.O = net.num_obs; fin.D = x.has_env ? cfg.num_dof : 0;
  fin.cfg = &cfg; fin.has_env = x.has_env;
  fin.actions_scaled = x.actions_scaled; fin.dof_targets = x.dof_targets; fin.reset = x.reset; fin.progress = x.progress;
  fin.potentials = x.potentials; fin.prev_potentials = x.prev_potentials; fin.root = x.root; fin.dof_state = x.dof_state;
  fin.t = &tab; fin.tw = &tab; fin.action_var = x.action_var; fin.rst_raw = 0; fin.trace = a.trace; fin.ti = 128;
  tcm::forward_persistent<true, false>(net, a, fin);
  if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[191] = clock64();       // kernel exit
}

// ------------------------------------------------------------------------------------------------ host side
static unsigned long long* g_trace = nullptr;   // debug timeline buffer (flyb_debug_set_trace), normally NULL
void tc_set_trace(void* p) { g_trace = (unsigned long long*)p; }
unsigned long long* tc_get_trace() { return g_trace; }
bool tc_supported(const FlybNetLayout* net) { return net->k0_pad == tcm::K0 && net->num_act <= tcm::APAD; }

// wpack = [weight stream: total_stages x 8 KB][inverse index: net->total x 4 int32]
size_t tc_wpack_bytes(const FlybNetLayout* net) {
  if (!tc_supported(net)) return 0;
  return (size_t)forward_plan(*net).total_stages * tcm::STAGE + (size_t)net->total * 16;
}

const int* tc_wpack_index(const FlybNetLayout* net, const void* wpack) {
  return (const int*)((const char*)wpack + (size_t)forward_plan(*net).total_stages * tcm::STAGE);
}

int tc_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, cudaStream_t st) {
  const PackPlan plan = forward_plan(*net);
  const int total = plan.total_stages * (tcm::STAGE / 4);
  int* inv = const_cast<int*>(tc_wpack_index(net, wpack));
  int rc = check_cuda(cudaMemsetAsync(inv, 0xFF, (size_t)net->total * 16, st), "cudaMemsetAsync(pack index)");
  if (rc != FLYB_OK) return rc;
  pack_index_kernel<<<std::min((total + 255) / 256, 2 * 148), 256, 0, st>>>(plan, inv);
  FLYB_LAUNCH_CHECK("pack_index_kernel");
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
                 mu_out ? (size_t)tcm::TM * net->num_act * 4 : 0, g_trace};
  const size_t smem = tcm::smem_bytes(net->num_obs, a.out_bytes);
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
  const size_t out = (size_t)tcm::TM * 3 * net->num_act * 4;
  return tcm::smem_bytes(net->num_obs, out) <= 227u * 1024u;
}

int tc_launch_act(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs, long long rows,
                  const float* action_var, const float* noise, uint64_t seed, uint64_t offset,
                  const uint64_t* offset_dev, float* action, float* logp, float* mu_out, float* obs_store,
                  const FlybEnvCfg* cfg, float* actions_scaled, float* dof_targets, int64_t* reset, int64_t* progress,
                  float* potentials, float* prev_potentials, float* root, float* dof_state, cudaStream_t st) {
  const size_t out = (size_t)tcm::TM * 3 * net->num_act * 4;
  tcm::FwdArgs a{obs, (const unsigned char*)wpack, params, rows, net->num_obs, net->num_act, out, g_trace};
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
  const size_t smem = tcm::smem_bytes(net->num_obs, out);
  const long long ntiles = (rows + tcm::TM - 1) / tcm::TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  int rc;
  if ((rc = prepare(policy_act_tc_kernel, smem)) != FLYB_OK) return rc;
  policy_act_tc_kernel<<<grid, tcm::THREADS, smem, st>>>(*net, a, x, c);
  FLYB_LAUNCH_CHECK("policy_act_tc_kernel");
  return FLYB_OK;
}

}  // namespace flyb
