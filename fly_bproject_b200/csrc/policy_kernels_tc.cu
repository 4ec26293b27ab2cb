// Kernel (b), tensor-core path: policy act (Net.pi forward + Gaussian sample / log-prob / clip + action scaling,
// DOF-target scatter and reset bookkeeping) and the plain Net.pi / Net.v forward, on tcgen05 / TMEM tiles fed by
// bulk TMA (mlp_tc.cuh).  Same results contract as the CUDA-core path in policy_kernels.cu (reference
// ppo.py:88-102, 210-220 and fly.py:626-660, 446-480); GEMMs are 3xTF32 with fp32 accumulation.
//
// Also here: flyb_pack_weights, which rewrites the flat fp32 parameters as the packed hi/lo weight stream the
// TMA producer reads (consumption order, 4 KB stages — layout in mlp_tc.cuh).
#include <algorithm>
#include <cstdlib>

#include "env_math.cuh"
#include "env_scan.cuh"
#include "mlp_tc.cuh"
#include "philox.cuh"

namespace flyb {

int num_sms();
int validate_net(const FlybNetLayout* net, const float* params);
unsigned long long* tc_get_trace();

namespace FLYB_TC_VNS {   // tcn (narrow shape) or tcw (wide shape, FLYB_TC_WIDE): see mlp_tc.cuh

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
    s.nstages = stages_for(N, ksteps);
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
  {
    auto addt = [&](int N, int ksteps, int w_off, int ldw, int n0, int n_rows, int k_cols) {
      add(N, ksteps, w_off, ldw, n0, 0, n_rows, k_cols);
      p.seg[p.nseg - 1].transposed = 1;
    };
    addt(H3, APAD / 8, net.off[6], H3, 0, H3, net.num_act);          // W4^T : dM1 = dZmu * W4       (k = action)
    addt(H2, H3 / 8, net.off[4], H2, 0, H2, H3);                     // W3pi^T: dH2 += dZm1 * W3pi    (k = m1 feature)
    addt(H2, H3 / 8, net.off[8], H2, 0, H2, H3);                     // W3v^T : dH2 += dZv1 * W3v
    addt(128, H2 / 8, net.off[2], H1, 0, 128, H2);                   // W2^T, h1 features [0,128)   (k = h2 feature)
    addt(128, H2 / 8, net.off[2], H1, 128, 128, H2);                 // W2^T, h1 features [128,256)
  }
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
    // a stage holds kps whole k-steps ([hi slab | lo slab] each); what is left of it (N = 48) stays zero
    const int kps = tcm::STAGE / (2 * s.N * 32);
    const int st_local = li / (tcm::STAGE / 4), within = li - st_local * (tcm::STAGE / 4);
    const int slab = within / slab_floats, e = within - slab * slab_floats;
    const int ks = st_local * kps + (slab >> 1), is_lo = slab & 1;
    const int kc = e / (s.N * 4), rem = e - kc * (s.N * 4);
    const int n = rem >> 2, kk = rem & 3;
    const int k = s.k0 + ks * 8 + kc * 4 + kk;
    float w = 0.0f;
    if ((slab >> 1) < kps && n < s.n_rows && k < s.k_cols)
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
    const int kps = tcm::STAGE / (2 * s.N * 32);
    const int st_local = li / (tcm::STAGE / 4), within = li - st_local * (tcm::STAGE / 4);
    const int slab = within / slab_floats, e = within - slab * slab_floats;
    if ((slab & 1) || (slab >> 1) >= kps) continue;   // lo slab: reached through delta; tail of the stage: unused
    const int ks = st_local * kps + (slab >> 1);
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
  static constexpr bool kBuildsObs = false;
  float* mu_out;
  float* value_out;
  int A;
  FLYB_D void build_obs(const tcm::Smem&, long long, int) {}
  FLYB_D void aux_warp(const tcm::Smem&, long long) {}
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
  static constexpr bool kBuildsObs = false;
  FLYB_D void build_obs(const tcm::Smem&, long long, int) {}
  FLYB_D void aux_warp(const tcm::Smem&, long long) {}
  int anneal_steps;        // variance annealing steps applied on the fly to action_var (fused rollout, ppo.py:236-237)
  float anneal_decay, anneal_floor;
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
      float var = __ldg(action_var + k);
      for (int j = 0; j < anneal_steps; ++j) var = fmaxf(anneal_floor, var - anneal_decay);   // iterated: same bits
      tab.std[k] = sqrtf(var);
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
  const float* obs_src;    // wide shape: the observation rows in global memory (there is no raw stage to copy from)
  FLYB_D void begin_tile(const tcm::Smem& s, long long row0, int r, int nrows) {
    rst_raw = (has_env && reset != nullptr && r < nrows) ? reset[row0 + r] : 0;
    if (obs_store) {   // rollout row all_obs[i] = obs (ppo.py:226)
      if constexpr (tcm::WIDE) {
        const float* src = obs_src + (size_t)row0 * O;
        float* dst = obs_store + (size_t)row0 * O;
        for (int i = threadIdx.x; i < nrows * O; i += tcm::EPI_THREADS) dst[i] = __ldg(src + i);
      } else {
        store_tile(obs_store + (size_t)row0 * O, s.obs, nrows * O);
        if (threadIdx.x == 0) tc::tma_store_commit();
      }
    }
  }
  FLYB_D void obs_consumed(const tcm::Smem& s) {
    if (!tcm::WIDE && obs_store && threadIdx.x == 0) tc::tma_store_wait_read<0>();   // the bulk store has read the staging buffer
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


#if !FLYB_TC_WIDE
// ------------------------------------------------------------------------------------------------ fused env step
// ONE launch per env step of a rollout: [get_obs + progress += 1 + get_reward of step i-1 (fly.py:676-681)] followed by
// [act + action scaling + DOF targets + Fly.reset bookkeeping of step i (ppo.py:213-220, fly.py:626-660, 446-480)] for
// the same 128 envs, in the CTA that owns them.  The epilogue warps build their tile's observation rows from the
// physics state straight into the shared-memory tile the MLP reads (while the weight stream is already in flight);
// the rows leave once, for the rollout ring.  Reset ids: the tile's 128-bit mask is published by a spare warp, and
// the spare warp of the tile that finishes last scans the bitmap — the MLP warps never wait for it.
struct StepTables {
  int foff[FORCE_VALS];     // float offset of gathered contact value rc inside one env's force block
  int aidx[FLYB_MAX_DOF];   // action k -> DOF
  float lo[FLYB_MAX_DOF], hi[FLYB_MAX_DOF];
  unsigned mask[2][4];      // [tile parity][32-env group]: reset flags of the tile
};

struct StepEnvArgs {
  const float* force;
  float* obs_out;        // nullable: env.obs_buf
  float* reward;
  float* reward_ring;    // nullable
  float* reward_sum;     // nullable
  float* up_vec;
  float* heading_vec;
  int* reset_ids;
  int* reset_count;
  ScanScratch* scratch;
  unsigned* tile_cnt;    // tagged per-tile reset counts behind the scratch's masks (env_scan.cuh)
};

constexpr int STEP_THREADS = tcm::THREADS + 32;   // + the mask / scan warp
constexpr int STEP_AUX_WARP = tcm::THREADS / 32;
constexpr int STEP_AUX_BAR_THREADS = tcm::EPI_THREADS + 32;
constexpr size_t STEP_ENV_STAGE_BYTES = (size_t)tcm::TM * (13 + FORCE_VALS + 2) * 4;   // root, force rows, two scalars

FLYB_D void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
FLYB_D float4 ld_global4(const float4* p) {   // coherent 128-bit load (the same kernel later rewrites reset rows)
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

struct StepFinal : ActFinal {
  static constexpr bool kBuildsObs = true;
  StepEnvArgs e;
  StepTables* st;
  bool tables_built;
  unsigned parity;
  // contact forces by bulk copy: the tile's 128 contiguous force blocks (128 x B x 12 bytes) land in the X | obs | out
  // region, free until the rows are built, and the 33 values an env needs are picked out of shared memory.  The gather
  // of 11-of-93 rows through the LSU touches ~940 of an env's 1116 bytes in DRAM anyway (64-byte access granularity)
  // and is bound by the SM's outstanding-miss limit; one sequential copy streams at the HBM rate.
  int force_bulk;          // 0: gather through the LSU
  unsigned force_phase;    // phase of s.obs_full
  bool first_tile;
  // back-pressure from the spare warp (persistent loop, more than one tile per CTA): it arrives here once it holds a
  // tile's masks in registers; the epilogue warps hand over the NEXT tile's masks (bar.arrive 3) only after that, so
  // named barrier 3 never collects two arrivals of the same warp and mask[][] is never overwritten unread
  uint64_t* aux_done;
  unsigned tiles_handed;   // tiles of this CTA whose masks the epilogue warps have handed to the spare warp

  int tb_;   // next env-phase trace slot (172..189), thread 0 of CTA 0
  FLYB_D void emark() {
    if (trace != nullptr && threadIdx.x == 0 && blockIdx.x == 0 && tb_ < 190) trace[tb_++] = clock64();
  }
  FLYB_D void build_obs(const tcm::Smem& s, long long row0, int nrows) {
    constexpr int TM = tcm::TM;
    emark();   // [0] env phase entered
    const FlybEnvCfg& c = *cfg;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int B = c.num_bodies;
    float* obs_s = s.obs;                              // [TM][O]
    float* root_s = s.out;                             // [TM][13]
    float* force_s = root_s + TM * 13;                 // [TM][33]
    int* cnt_s = reinterpret_cast<int*>(force_s + TM * FORCE_VALS);   // legs | at_limit << 8 | abdomen_hit << 16
    float* sq_s = reinterpret_cast<float*>(cnt_s + TM);
    tcm::epi_bar_sync();   // the previous tile's pi() is done with s.out; its ring store has drained s.obs (obs_consumed)
    emark();   // [1] all eight warps are here
    if (!tables_built) {
      if (tid < FORCE_VALS) {
        const int r = tid / 3;
        const int body = (r < FLYB_NUM_ABDOMEN) ? c.abdomen_local[r] : c.legtip_local[r - FLYB_NUM_ABDOMEN];
        st->foff[tid] = body * 3 + (tid - r * 3);
      }
      if (tid >= 64 && tid < 64 + A) {
        const int d = c.action_idx[tid - 64];
        st->aidx[tid - 64] = d;
        st->lo[tid - 64] = c.lower[d];
        st->hi[tid - 64] = c.upper[d];
      }
      tables_built = true;
      tcm::epi_bar_sync();
    }
    emark();   // [2] tables built
    // ---- every global load of the tile is issued before the first dependent shared-memory store
    const int en = tid >> 1, qd = tid & 1;             // two threads per env for the gathers
    const bool e_ok = en < nrows;
    const long long env = row0 + en;
    const unsigned fbytes = (unsigned)nrows * (unsigned)B * 12u;
    const bool bulk = force_bulk != 0 && (fbytes & 15u) == 0u;
    if (bulk && !first_tile && tid == 0) {             // (the first tile's copy was issued at kernel entry)
      tc::fence_proxy_async();
      tc::mbar_arrive_expect_tx(s.obs_full, fbytes);
      tc::tma_load_1d(s.xhi, reinterpret_cast<const char*>(e.force) + (size_t)row0 * B * 12, fbytes, s.obs_full);
    }
    first_tile = false;
    const int nroot = nrows * 13;
    const float4* root4 = reinterpret_cast<const float4*>(root + (size_t)row0 * 13);   // row0 % 128 == 0: 16-byte aligned
    float4 r4[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int idx = tid + tcm::EPI_THREADS * u;
      r4[u] = (idx * 4 + 3 < nroot) ? ld_global4(root4 + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float pot_in[2] = {0.f, 0.f};
    long long prog_in[2] = {0, 0}, reset_in[2] = {0, 0};
    if (tid < 64) {
#pragma unroll
      for (int p = 0; p < 2; ++p)
        if (p * 64 + tid < nrows) {
          pot_in[p] = potentials[row0 + p * 64 + tid];
          prog_in[p] = progress[row0 + p * 64 + tid];
          reset_in[p] = reset[row0 + p * 64 + tid];
        }
    }
    emark();   // [3] root / scalar loads issued
    constexpr int FITER = (FORCE_VALS + 1) / 2;   // 17
    float fv[FITER];
    if (!bulk) {
      const float* fe = e.force + (size_t)env * B * 3;
#pragma unroll
      for (int it = 0; it < FITER; ++it) {
        const int rc = qd + 2 * it;
        fv[it] = (e_ok && rc < FORCE_VALS) ? __ldg(fe + st->foff[rc]) : 0.0f;
      }
    }
    // DOF pairs / commanded actions: the tile's pairs as ONE flat array, element f = (env f / A, action f % A) — a warp
    // reads 32 consecutive (or, with a partial actuation, nearly consecutive) pairs: coalesced, a handful of lines per
    // instruction instead of one line per lane
    constexpr int DU = 10;                        // ceil(128 x 20 / 256): A <= 20 on this path (O = 19 + 3A <= 80)
    float2 pv[DU];
    float cmd[DU];
    const int npairs = nrows * A;
    const int el_step = tcm::EPI_THREADS / A, k_step = tcm::EPI_THREADS % A;
    {
      const float2* de = reinterpret_cast<const float2*>(dof_state) + (size_t)row0 * D;
      const float* ae = actions_scaled + (size_t)row0 * A;
      int el = tid / A, k = tid - el * A;
#pragma unroll
      for (int u = 0; u < DU; ++u) {
        const int f = tid + tcm::EPI_THREADS * u;
        pv[u] = make_float2(0.f, 0.f);
        cmd[u] = 0.0f;
        if (f < npairs) {
          pv[u] = de[el * D + st->aidx[k]];
          cmd[u] = ae[f];
        }
        el += el_step;
        k += k_step;
        if (k >= A) { k -= A; ++el; }
      }
    }
    emark();   // [4] every load issued
    if (bulk) {   // the force blocks have landed: pick this env's 33 values, then hand the region back
      tc::mbar_wait(s.obs_full, force_phase);
      force_phase ^= 1u;
      const float* fe = s.xhi + (size_t)en * B * 3;     // block stride 279 words: odd, so conflict-free across envs
#pragma unroll
      for (int it = 0; it < FITER; ++it) {
        const int rc = qd + 2 * it;
        fv[it] = (e_ok && rc < FORCE_VALS) ? fe[st->foff[rc]] : 0.0f;
      }
      tcm::epi_bar_sync();
    }
    // ---- stage: obs[12 : 12+3A] (fly.py:795, 801, 802), root rows, gathered contact rows
    {
      int el = tid / A, k = tid - el * A;
#pragma unroll
      for (int u = 0; u < DU; ++u) {
        const int f = tid + tcm::EPI_THREADS * u;
        if (f < npairs) {
          float* oe = obs_s + el * O + 12;
          oe[k] = unscale(pv[u].x, st->lo[k], st->hi[k]);
          oe[A + k] = pv[u].y * c.dof_vel_scale;
          oe[2 * A + k] = cmd[u];
        }
        el += el_step;
        k += k_step;
        if (k >= A) { k -= A; ++el; }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int idx = tid + tcm::EPI_THREADS * u;
      if (idx * 4 + 3 < nroot) reinterpret_cast<float4*>(root_s)[idx] = r4[u];
    }
    for (int i = (nroot & ~3) + tid; i < nroot; i += tcm::EPI_THREADS) root_s[i] = root[(size_t)row0 * 13 + i];
#pragma unroll
    for (int it = 0; it < FITER; ++it) {
      const int rc = qd + 2 * it;
      if (e_ok && rc < FORCE_VALS) force_s[en * FORCE_VALS + rc] = fv[it];
    }
    tcm::epi_bar_sync();
    emark();   // [5] loads consumed, tile staged

    // ---- per-env math: four roles (warp pairs), each over the tile's two 64-env halves.  role 0: potential, heading /
    //      up projections, later reward + termination; 1: local velocities; 2: Euler angles; 3: contact flags, sums
    const int role = warp >> 1;
    const int me0 = ((warp & 1) << 5) | lane;
    // (the two passes stay rolled: this code runs once per launch, every instruction-cache line it spans is a miss)
    float k_hproj[2], k_pot[2], k_h[2], k_qz[2], k_qw[2];
    Vec3 k_up[2], k_head[2];
#pragma unroll 1
    for (int p = 0; p < 2; ++p) {
      const int me = p * 64 + me0;
      const bool me_ok = me < nrows;
      float r[13];
#pragma unroll
      for (int j = 0; j < 13; ++j) r[j] = me_ok ? root_s[me * 13 + j] : (j == 6 ? 1.0f : 0.0f);
      float* o = obs_s + me * O;
      const Quat tq = env_torso_quat(r, c);
      EnvObsHead h;
      h.heading_proj = 0.0f; h.potential = 0.0f;
      if (role == 0) {
        env_head_heading(r, c, tq, h);
        if (me_ok) { o[10] = h.up_proj; o[11] = h.heading_proj; }
        if (p == 0) {
          k_hproj[0] = h.heading_proj; k_pot[0] = h.potential; k_h[0] = r[2]; k_qz[0] = r[5]; k_qw[0] = r[6];
          k_up[0] = h.up_vec; k_head[0] = h.heading_vec;
        } else {
          k_hproj[1] = h.heading_proj; k_pot[1] = h.potential; k_h[1] = r[2]; k_qz[1] = r[5]; k_qw[1] = r[6];
          k_up[1] = h.up_vec; k_head[1] = h.heading_vec;
        }
      } else if (role == 1) {
        env_head_vel(r, tq, h);
        if (me_ok) {
          o[0] = h.height;
          o[1] = h.vel_loc.x; o[2] = h.vel_loc.y; o[3] = h.vel_loc.z;
          o[4] = h.angvel_loc.x; o[5] = h.angvel_loc.y; o[6] = h.angvel_loc.z;
        }
      } else if (role == 2) {
        env_head_euler(r, c, tq, h);
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
          o[13 + 3 * A + b] = touch ? 1.0f : 0.0f;    // fly.py:797, 803
          legs += touch ? 1 : 0;
        }
        float actions_sq_sum = 0.0f;
        const float* cm = o + 12 + 2 * A;
        for (int k = 0; k < A; ++k) {
          const float av = cm[k];
          actions_sq_sum += av * av;                    // fly.py:731
          at_limit += (av > st->hi[k] * 0.9f) ? 1 : 0;  // fly.py:736-737
          at_limit += (av < st->lo[k] * 0.9f) ? 1 : 0;
        }
        cnt_s[me] = legs | (at_limit << 8) | ((abdomen > 0.0f ? 1 : 0) << 16);
        sq_s[me] = actions_sq_sum;
      }
    }
    tcm::epi_bar_sync();
    emark();   // [6] per-env math done
    if (role == 0) {
      float my_reward = 0.0f;
#pragma unroll 1
      for (int p = 0; p < 2; ++p) {
        const int me = p * 64 + me0;
        const bool me_ok = me < nrows;
        bool flag = false;
        // (static register indexing: select after reading both)
        const float s_h = p ? k_h[1] : k_h[0], s_hproj = p ? k_hproj[1] : k_hproj[0], s_qz = p ? k_qz[1] : k_qz[0];
        const float s_qw = p ? k_qw[1] : k_qw[0], s_pot = p ? k_pot[1] : k_pot[0], s_potin = p ? pot_in[1] : pot_in[0];
        const long long s_prog = p ? prog_in[1] : prog_in[0], s_rst = p ? reset_in[1] : reset_in[0];
        const Vec3 s_up = p ? k_up[1] : k_up[0], s_head = p ? k_head[1] : k_head[0];
        if (me_ok) {
          const long long gi = row0 + me;
          const int packed = cnt_s[me];
          EnvRewardIn in;
          in.actions_sq_sum = sq_s[me];
          in.electricity = 0.0f;
          in.at_limit = (packed >> 8) & 0xff;
          in.height = s_h;
          in.heading_proj = s_hproj;
          in.qz = s_qz;
          in.qw = s_qw;
          in.prev_potential = s_potin;      // fly.py:780
          in.potential = s_pot;
          in.legs_touching = packed & 0xff;
          in.abdomen_hit = (packed >> 16) != 0;
          long long prog = s_prog + 1;      // fly.py:678
          if (prog == 0) prog = 1;          // fly.py:415-416
          in.progress = prog;
          in.reset_prev = s_rst;
          float rew;
          long long rst;
          env_reward_reset(in, c, &rew, &rst);
          potentials[gi] = s_pot;
          prev_potentials[gi] = s_potin;
          progress[gi] = prog;
          reset[gi] = rst;                  // read back by this tile's act half (begin_tile), behind the barrier below
          e.reward[gi] = rew;
          if (e.reward_ring) e.reward_ring[gi] = rew;
          my_reward += rew;
          e.up_vec[gi * 3 + 0] = s_up.x; e.up_vec[gi * 3 + 1] = s_up.y; e.up_vec[gi * 3 + 2] = s_up.z;
          e.heading_vec[gi * 3 + 0] = s_head.x; e.heading_vec[gi * 3 + 1] = s_head.y;
          e.heading_vec[gi * 3 + 2] = s_head.z;
          flag = rst != 0;
        }
        const unsigned ballot = __ballot_sync(0xffffffffu, flag);
        if (lane == 0) st->mask[parity][p * 2 + (warp & 1)] = ballot;
      }
      if (e.reward_sum) {   // score accumulator (ppo.py:233), one atomic per warp
        const float ws = warp_sum(my_reward);
        if (lane == 0) atomicAdd(e.reward_sum, ws);
      }
    }
    tc::fence_proxy_async();   // the tile's rows in s.obs -> visible to the bulk stores of begin_tile
    // the spare warp has taken the previous tile's masks (its arrival number `tiles_handed - 1`, phases alternate)
    if (tiles_handed > 0u && threadIdx.x == 0) tc::mbar_wait(aux_done, (tiles_handed - 1u) & 1u);
    ++tiles_handed;
    tcm::epi_bar_sync();
    asm volatile("bar.arrive 3, %0;" ::"n"(STEP_AUX_BAR_THREADS) : "memory");   // masks are in shared memory: wake the spare warp
    parity ^= 1u;
    // ---- the CTA's NEXT tile (persistent loop, N > 128 x CTAs): pull its state lines into L2 now, so that the DRAM
    //      latency of its gathers is paid under this tile's MMAs.  Same addresses, same thread mapping as above.
    {
      const long long next0 = row0 + (long long)gridDim.x * TM;
      const long long nenv = c.num_envs;
      if (next0 + en < nenv) {
        const long long nx = next0 + en;
        const float* fe = e.force + (size_t)nx * B * 3;
#pragma unroll
        for (int it = 0; it < FITER; ++it) {
          const int rc = qd + 2 * it;
          if (rc < FORCE_VALS && (rc % 3) == 0) prefetch_l2(fe + st->foff[rc]);   // one touch per gathered row
        }
        if (qd == 0) {
          const char* dp = reinterpret_cast<const char*>(reinterpret_cast<const float2*>(dof_state) + (size_t)nx * D);
          for (int b = 0; b < D * 8; b += 128) prefetch_l2(dp + b);
          prefetch_l2(root + (size_t)nx * 13);
        } else {
          const char* ap = reinterpret_cast<const char*>(actions_scaled + (size_t)nx * A);
          for (int b = 0; b < A * 4; b += 128) prefetch_l2(ap + b);
          prefetch_l2(potentials + nx);
          prefetch_l2(progress + nx);
          prefetch_l2(reset + nx);
        }
      }
    }
    emark();   // [7] reward + masks done, rows visible to the bulk stores
  }

  FLYB_D void begin_tile(const tcm::Smem& s, long long row0, int r, int nrows) {
    rst_raw = (r < nrows) ? reset[row0 + r] : 0;
    if (obs_store) store_tile(obs_store + (size_t)row0 * O, s.obs, nrows * O);      // rollout row (ppo.py:226, 228)
    if (e.obs_out) store_tile(e.obs_out + (size_t)row0 * O, s.obs, nrows * O);      // env.obs_buf
    if ((obs_store || e.obs_out) && threadIdx.x == 0) tc::tma_store_commit();
  }
  FLYB_D void obs_consumed(const tcm::Smem& s) {
    if ((obs_store || e.obs_out) && threadIdx.x == 0) tc::tma_store_wait_read<0>();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tc::mbar_arrive(s.obs_empty);
  }

  // the spare warp: the tile's share of the reset-id compaction (env_scan.cuh: every tile writes its own ids once the
  // tiles in front have published their counts); the last tile to get there closes the launch's ticket
  FLYB_D void aux_warp(const tcm::Smem&, long long ntiles) {
    if ((threadIdx.x >> 5) != STEP_AUX_WARP) return;
    const int lane = threadIdx.x & 31;
    const unsigned tag = ld_relaxed_u32(&e.scratch->ticket) + 1u;   // bumped only after every tile of this launch is through
    unsigned par = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, par ^= 1u) {
      asm volatile("bar.sync 3, %0;" ::"n"(STEP_AUX_BAR_THREADS) : "memory");
      const unsigned m[4] = {st->mask[par][0], st->mask[par][1], st->mask[par][2], st->mask[par][3]};
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(aux_done);   // masks taken: the epilogue warps may hand over the next tile's
      compact_tile_warp(e.tile_cnt, tag, tile, ntiles, m, e.reset_ids, e.reset_count);
      if (lane == 0) {
        __threadfence();
        if (atomicAdd(&e.scratch->done, 1u) == (unsigned)(ntiles - 1)) {
          e.scratch->done = 0u;        // re-armed for the next launch (of this kernel or of post_physics_kernel)
          e.scratch->ticket = tag;
        }
      }
    }
  }
};
#endif   // !FLYB_TC_WIDE

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
  fin.action = x.action; fin.logp = x.logp; fin.mu_out = x.mu_out; fin.obs_store = x.obs_store; fin.obs_src = a.obs;
  fin.A = A; fin.O = net.num_obs; fin.D = x.has_env ? cfg.num_dof : 0;
  fin.cfg = &cfg; fin.has_env = x.has_env;
  fin.actions_scaled = x.actions_scaled; fin.dof_targets = x.dof_targets; fin.reset = x.reset; fin.progress = x.progress;
  fin.potentials = x.potentials; fin.prev_potentials = x.prev_potentials; fin.root = x.root; fin.dof_state = x.dof_state;
  fin.t = &tab; fin.tw = &tab; fin.action_var = x.action_var; fin.rst_raw = 0; fin.trace = a.trace; fin.ti = 128;
  fin.anneal_steps = 0; fin.anneal_decay = 0.0f; fin.anneal_floor = 0.0f;
  tcm::forward_persistent<true, false>(net, a, fin);
  if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[191] = clock64();       // kernel exit
}


#if !FLYB_TC_WIDE
struct StepTcArgs {
  ActTcArgs act;
  StepEnvArgs env;
  int anneal_steps;
  float anneal_decay, anneal_floor;
};

__global__ void __launch_bounds__(STEP_THREADS, 1)
env_step_tc_kernel(const __grid_constant__ FlybNetLayout net, const __grid_constant__ tcm::FwdArgs a,
                   const __grid_constant__ StepTcArgs y, const __grid_constant__ FlybEnvCfg cfg) {
  __shared__ ActTables tab;
  __shared__ StepTables stab;
  __shared__ __align__(8) uint64_t aux_done_bar;
  if (threadIdx.x == 0) {   // every other thread passes forward_persistent's opening __syncthreads() before using it
    tc::mbar_init(&aux_done_bar, 1);
    tc::fence_mbar_init();
  }
  const ActTcArgs& x = y.act;
  StepFinal fin;
  fin.aux_done = &aux_done_bar; fin.tiles_handed = 0u;
  fin.noise = x.noise; fin.seed = x.seed;
  fin.draw = x.offset + (x.offset_dev ? __ldg(x.offset_dev) : 0ull);
  fin.action = x.action; fin.logp = x.logp; fin.mu_out = x.mu_out; fin.obs_store = x.obs_store;
  fin.A = net.num_act; fin.O = net.num_obs; fin.D = cfg.num_dof;
  fin.cfg = &cfg; fin.has_env = 1;
  fin.actions_scaled = x.actions_scaled; fin.dof_targets = x.dof_targets; fin.reset = x.reset; fin.progress = x.progress;
  fin.potentials = x.potentials; fin.prev_potentials = x.prev_potentials; fin.root = x.root; fin.dof_state = x.dof_state;
  fin.t = &tab; fin.tw = &tab; fin.action_var = x.action_var; fin.rst_raw = 0; fin.trace = a.trace; fin.ti = 128;
  fin.anneal_steps = y.anneal_steps; fin.anneal_decay = y.anneal_decay; fin.anneal_floor = y.anneal_floor;
  fin.e = y.env; fin.st = &stab; fin.tables_built = false; fin.parity = 0u; fin.tb_ = 172;
  fin.force_bulk = a.pre_src != nullptr ? 1 : 0; fin.force_phase = 0u; fin.first_tile = true;
  if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[190] = clock64();   // kernel entry
  // debug: every CTA's entry / exit on the GLOBAL timer (ns), slots [192 + 2 b, 193 + 2 b]: launch skew, slow CTAs
  unsigned long long t_in = 0;
  if (a.trace != nullptr && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_in));
  tcm::forward_persistent<true, false>(net, a, fin);
  if (a.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) a.trace[191] = clock64();       // kernel exit
  if (a.trace != nullptr && threadIdx.x == 0 && blockIdx.x < 148) {
    unsigned long long t_out;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_out));
    a.trace[192 + 2 * blockIdx.x] = t_in;
    a.trace[193 + 2 * blockIdx.x] = t_out;
  }
}
#endif   // !FLYB_TC_WIDE

// ------------------------------------------------------------------------------------------------ host side
#define g_trace (::flyb::tc_get_trace())
// narrow: exactly the K0 = 80 layout (the PPO kernels' layer-1 accumulators assume it); wide: anything up to K0 = 160
bool tc_supported(const FlybNetLayout* net) {
  return (tcm::WIDE ? net->k0_pad <= tcm::K0 : net->k0_pad == tcm::K0) && net->num_act <= tcm::APAD;
}

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
  if (tcm::WIDE && out > 2 * tcm::X_PLANE_BYTES) return false;   // the staging aliases the X planes
  return tcm::smem_bytes(net->num_obs, out) + 3072 <= 227u * 1024u;
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


#if !FLYB_TC_WIDE
// fused env step (flyb_env_step_fused): shapes the tensor-core act path covers, with the env staging next to the ring
bool tc_step_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg) {
  if (!tc_supported(net) || cfg == nullptr || net->num_act > 20) return false;
  const size_t out = std::max((size_t)tcm::TM * 3 * net->num_act * 4, STEP_ENV_STAGE_BYTES);
  return tcm::smem_bytes(net->num_obs, out) <= 227u * 1024u - 3072u;   // static tables of the kernel
}

int tc_launch_step(const FlybNetLayout* net, const float* params, const void* wpack, const FlybEnvCfg* cfg,
                   const FlybStepIO* io, cudaStream_t st) {
  const size_t out = std::max((size_t)tcm::TM * 3 * net->num_act * 4, STEP_ENV_STAGE_BYTES);
  const long long rows = cfg->num_envs;
  tcm::FwdArgs a{nullptr, (const unsigned char*)wpack, params, rows, net->num_obs, net->num_act, out, g_trace};
  static const bool no_bulk = [] { const char* ev = getenv("FLYB_STEP_FORCE_BULK"); return ev && ev[0] == '0'; }();
  const size_t block = (size_t)cfg->num_bodies * 12;
  if (!no_bulk && (size_t)tcm::TM * block <= tcm::build_region_bytes(net->num_obs, out) && aligned16(io->force)) {
    a.pre_src = io->force;
    a.pre_row_bytes = (unsigned)block;
  }
  StepTcArgs y{};
  ActTcArgs& x = y.act;
  x.action_var = io->action_var; x.noise = io->noise; x.seed = io->seed; x.offset = io->offset;
  x.offset_dev = (const unsigned long long*)io->offset_dev;
  x.action = io->action; x.logp = io->logp; x.mu_out = nullptr; x.obs_store = io->obs_ring;
  x.has_env = 1;
  x.actions_scaled = io->actions_scaled; x.dof_targets = io->dof_targets; x.reset = (long long*)io->reset;
  x.progress = (long long*)io->progress; x.potentials = io->potentials; x.prev_potentials = io->prev_potentials;
  x.root = io->root; x.dof_state = io->dof_state;
  y.env.force = io->force; y.env.obs_out = io->obs; y.env.reward = io->reward; y.env.reward_ring = io->reward_ring;
  y.env.reward_sum = io->reward_sum; y.env.up_vec = io->up_vec; y.env.heading_vec = io->heading_vec;
  y.env.reset_ids = io->reset_ids; y.env.reset_count = io->reset_count; y.env.scratch = (ScanScratch*)io->scratch;
  y.env.tile_cnt = scan_tile_counts(y.env.scratch, cfg->num_envs);
  y.anneal_steps = io->anneal_steps; y.anneal_decay = io->anneal_decay; y.anneal_floor = io->anneal_floor;
  const size_t smem = tcm::smem_bytes(net->num_obs, out);
  const long long ntiles = (rows + tcm::TM - 1) / tcm::TM;
  const int grid = (int)std::min<long long>(ntiles, (long long)num_sms());
  int rc;
  if ((rc = prepare(env_step_tc_kernel, smem)) != FLYB_OK) return rc;
  env_step_tc_kernel<<<grid, STEP_THREADS, smem, st>>>(*net, a, y, *cfg);
  FLYB_LAUNCH_CHECK("env_step_tc_kernel");
  return FLYB_OK;
}

#endif   // !FLYB_TC_WIDE

}  // namespace FLYB_TC_VNS

#if !FLYB_TC_WIDE
static unsigned long long* g_trace_buf = nullptr;   // debug timeline buffer (flyb_debug_set_trace), normally NULL
void tc_set_trace(void* p) { g_trace_buf = (unsigned long long*)p; }
unsigned long long* tc_get_trace() { return g_trace_buf; }
#endif

}  // namespace flyb
