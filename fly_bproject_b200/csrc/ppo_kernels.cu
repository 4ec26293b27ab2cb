// Kernel (d): one optimiser step of PPO.update — forward, clipped-surrogate + SmoothL1 loss, backward,
// global-norm clip and Adam — as the exact-fp32 CUDA-core path.
//
// Replaces reference ppo.py:183-199 (~100 autograd launches per optimiser step, the trunk evaluated twice,
// a Cholesky + MultivariateNormal per minibatch).  Structure:
//   ppo_fwd_bwd_kernel   per 64-row tile, everything in shared memory: forward of both heads on ONE trunk
//                        evaluation, loss terms, dL/dmu and dL/dv, then the backward chain through the
//                        activations (dZ = (dY * W) .* ELU').  Writes the layer inputs X_l and the dZ_l to HBM.
//   ppo_dw_kernel        dW_l = dZ_l^T * X_l and db_l = colsum(dZ_l) for all six layers in one launch,
//                        split over row ranges; each split writes its own partial (deterministic, no atomics).
//   ppo_reduce_kernel    grad = sum of partials; loss = mean(-surr) + mean(smooth_l1).
//   [data-parallel runs all-reduce `grad` here]
//   ppo_norm_kernel / ppo_adam_kernel   clip_grad_norm_ (ppo.py:198) + torch.optim.Adam.step (ppo.py:199).
#include <algorithm>
#include <cstdlib>

#include "adam.cuh"
#include "peer.cuh"
#include "tc_common.cuh"
#include "tc_dispatch.cuh"
#include "mlp_simt.cuh"

namespace flyb {

int num_sms();
int validate_net(const FlybNetLayout* net, const float* params);
// flyb_ppo_step: the gradient reduction also takes the squared norm and bumps Adam's step counter, so that Adam can
// follow without a norm kernel.  in: step;  out: sq_part [sq_parts]
#define FLYB_STEP_FUSE_DEFINED
struct StepFuse {
  long long* step;
  const float* sq_part;
  int sq_parts;
  // when `tail` is set the weight-gradient kernel itself finishes the step (reduction + clip + Adam behind grid barriers)
  int tail;
  float *params, *exp_avg, *exp_avg_sq, *grad_norm_out;
  double lr, beta1, beta2, eps, max_norm;
  int done;   // out: 1 = the step was completed by the tail (no Adam launch needed); 2 = the reduction did the
              // fused exchange (grad = sum over ranks, sq_part filled): clip + Adam are left; 3 = the reduction also ran
              // clip + Adam (no launch left)
  float grad_scale;   // 1 / world when the tail also does the data-parallel exchange (peers != NULL), else 1
  int exchange;   // in: with peers and no tail, the reducing kernel does the exchange itself (flyb_ppo_step_dp)
};
constexpr int DW_SPLITS_MAX = 64;
constexpr int NORM_BLOCKS = 256;

struct PpoWs {   // device pointers into the caller's workspace
  float *h1, *h2, *m1, *v1;            // layer inputs saved by the forward
  float *dzh1, *dzh2, *dzm1, *dzv1, *dzmu, *dzval;
  float* loss_part;                    // [ntiles][2]
  float* dw_part;                      // [splits][net.total]
  float* norm_part;                    // [NORM_BLOCKS]
  int splits;
  long long ntiles;
  size_t bytes;
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static PpoWs carve_ws(const FlybNetLayout& net, long long rows, void* base) {
  PpoWs w{};
  char* p = (char*)base;
  size_t off = 0;
  auto take = [&](size_t nfloats) {
    float* r = (float*)(p + off);
    off += align_up(nfloats * sizeof(float), 256);
    return r;
  };
  const size_t R = (size_t)rows;
  w.norm_part = take(NORM_BLOCKS);   // first slot: flyb_ppo_apply finds it without knowing `rows`
  w.ntiles = (rows + MLP_TM - 1) / MLP_TM;
  // enough splits to fill the machine (~21 output tiles per split), each split a multiple of 16 rows
  w.splits = (int)std::max<long long>(1, std::min<long long>(DW_SPLITS_MAX, rows / 512));
  w.h1 = take(R * MLP_H1); w.h2 = take(R * MLP_H2); w.m1 = take(R * MLP_H3); w.v1 = take(R * MLP_H3);
  w.dzh1 = take(R * MLP_H1); w.dzh2 = take(R * MLP_H2); w.dzm1 = take(R * MLP_H3); w.dzv1 = take(R * MLP_H3);
  w.dzmu = take(R * net.num_act); w.dzval = take(R);
  w.loss_part = take((size_t)w.ntiles * 2);
  w.dw_part = take((size_t)w.splits * net.total);
  w.bytes = off;
  return w;
}

struct GradArgs {
  const float* obs;
  const float* action;
  const float* old_logp;
  const float* target;
  const float* advantage;
  const float* action_var;
  long long rows;
  float clip_lo, clip_hi;   // 1 - clip, 1 + clip rounded to fp32 as torch.clamp does with Python scalars
  float inv_rows;
};

// ------------------------------------------------------------------------------------------ forward + backward chain
template <int APAD>
__global__ void __launch_bounds__(MLP_THREADS, 1)
ppo_fwd_bwd_kernel(const __grid_constant__ FlybNetLayout net, const float* __restrict__ params,
                   const __grid_constant__ GradArgs a, const __grid_constant__ PpoWs w) {
  extern __shared__ __align__(16) float smem[];
  const MlpSmem s = mlp_carve(smem, net.k0_pad);
  __shared__ float s_std[FLYB_MAX_DOF];
  __shared__ float s_half_log_det;
  __shared__ float s_red[2][MLP_THREADS / 32];
  const int A = net.num_act, O = net.num_obs;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cg = lane, rg = warp;
  const float* p = params;

  if (threadIdx.x < A) s_std[threadIdx.x] = sqrtf(__ldg(a.action_var + threadIdx.x));
  __syncthreads();
  if (threadIdx.x == 0) {
    float hl = 0.0f;
    for (int k = 0; k < A; ++k) hl += logf(s_std[k]);
    s_half_log_det = hl;
  }
  const float a_log_2pi = (float)((double)A * 1.8378770664093453);

  for (long long tile = blockIdx.x; tile < w.ntiles; tile += gridDim.x) {
    const long long row0 = tile * MLP_TM;
    const int nrows = (int)min((long long)MLP_TM, a.rows - row0);
    mlp_load_obs_tile(s, a.obs, row0, nrows, O, net.k0_pad);
    MlpSave save;
    save.h1 = w.h1 + (size_t)row0 * MLP_H1; save.h2 = w.h2 + (size_t)row0 * MLP_H2;
    save.m1 = w.m1 + (size_t)row0 * MLP_H3; save.v1 = w.v1 + (size_t)row0 * MLP_H3;
    save.rows = nrows;
    mlp_forward_tile<true, true, APAD>(s, net, params, save);

    // ---- loss terms and dL/d(pre-activation) of the two heads: one warp per row (ppo.py:184-197)
    float surr_sum = 0.0f, sl1_sum = 0.0f;
    for (int r = warp; r < MLP_TM; r += MLP_THREADS / 32) {
      float* mu = s.buf0 + r * MLP_LD0 + MLP_COL_MU;
      float* val = s.buf0 + r * MLP_LD0 + MLP_COL_VAL;
      if (r >= nrows) {   // padded rows contribute nothing downstream
        for (int k = lane; k < APAD; k += 32) mu[k] = 0.0f;
        if (lane == 0) *val = 0.0f;
        continue;
      }
      const long long row = row0 + r;
      float m = 0.0f;
      for (int k = lane; k < A; k += 32) {
        const float diff = (__ldg(a.action + (size_t)row * A + k) - mu[k]) / s_std[k];
        m += diff * diff;
      }
      m = warp_sum(m);
      const float logp = -0.5f * (a_log_2pi + m) - s_half_log_det;
      const float ratio = expf(logp - __ldg(a.old_logp + row));
      const float adv = __ldg(a.advantage + row);
      const float s1 = ratio * adv;
      const float s2 = fminf(fmaxf(ratio, a.clip_lo), a.clip_hi) * adv;
      const bool in_range = ratio >= a.clip_lo && ratio <= a.clip_hi;
      // d min(s1, s2) / d ratio with torch.minimum's tie rule (half to each side)
      float g;
      if (s1 < s2) g = adv;
      else if (s1 > s2) g = in_range ? adv : 0.0f;
      else g = 0.5f * adv + (in_range ? 0.5f * adv : 0.0f);
      const float dlogp = -a.inv_rows * g * ratio;   // d mean(-surr) / d logp
      for (int k = lane; k < A; k += 32) {
        const float mu_k = mu[k];
        const float dmu = dlogp * ((__ldg(a.action + (size_t)row * A + k) - mu_k) / (s_std[k] * s_std[k]));
        const float dz = dmu * elu_grad_from_out(mu_k);   // trailing ELU on mu (ppo.py:30)
        mu[k] = dz;
        w.dzmu[(size_t)row * A + k] = dz;
      }
      const float z = *val - __ldg(a.target + row);
      const float az = fabsf(z);
      const float dval = a.inv_rows * (az < 1.0f ? z : (z > 0.0f ? 1.0f : -1.0f));   // smooth_l1, beta = 1
      __syncwarp();
      if (lane == 0) {
        surr_sum += fminf(s1, s2);
        sl1_sum += az < 1.0f ? 0.5f * z * z : az - 0.5f;
        *val = dval;
        w.dzval[row] = dval;
      }
    }
    if (lane == 0) { s_red[0][warp] = surr_sum; s_red[1][warp] = sl1_sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float x = 0.0f, y = 0.0f;
      for (int i = 0; i < MLP_THREADS / 32; ++i) { x += s_red[0][i]; y += s_red[1][i]; }
      w.loss_part[tile * 2 + 0] = x;
      w.loss_part[tile * 2 + 1] = y;
    }

    // ---- backward chain.  buf0: m1 [0,64) | dZmu [64,64+APAD) | v1 [128,192) | dZval [192];  buf1: h2
    {  // dZm1 = (dZmu * Wm2) .* ELU'(m1), in place over m1
      float acc[8][2];
#pragma unroll
      for (int r = 0; r < 8; ++r) acc[r][0] = acc[r][1] = 0.0f;
      mlp_layer_nt_acc<MLP_H3>(acc, p + net.off[6], APAD, A, s.buf0 + MLP_COL_MU, MLP_LD0, s.wst);
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const int rr = rg * 8 + r, n = cg + 32 * j;
          float* y = s.buf0 + rr * MLP_LD0 + MLP_COL_M1 + n;
          const float dz = acc[r][j] * elu_grad_from_out(*y);
          *y = dz;
          if (rr < nrows) w.dzm1[(size_t)(row0 + rr) * MLP_H3 + n] = dz;
        }
    }
    {  // dZv1 = (dZval * Wv2) .* ELU'(v1), in place over v1
      const float w0 = __ldg(p + net.off[10] + cg), w1 = __ldg(p + net.off[10] + cg + 32);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int rr = rg * 8 + r;
        const float dv = s.buf0[rr * MLP_LD0 + MLP_COL_VAL];
        float* y0 = s.buf0 + rr * MLP_LD0 + MLP_COL_V1 + cg;
        const float d0 = dv * w0 * elu_grad_from_out(y0[0]);
        const float d1 = dv * w1 * elu_grad_from_out(y0[32]);
        y0[0] = d0;
        y0[32] = d1;
        if (rr < nrows) {
          w.dzv1[(size_t)(row0 + rr) * MLP_H3 + cg] = d0;
          w.dzv1[(size_t)(row0 + rr) * MLP_H3 + cg + 32] = d1;
        }
      }
    }
    __syncthreads();
    {  // dZh2 = (dZm1 * Wm1 + dZv1 * Wv1) .* ELU'(h2), in place over h2 (buf1)
      float acc[8][4];
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[r][j] = 0.0f;
      mlp_layer_nt_acc<MLP_H2>(acc, p + net.off[4], MLP_H3, MLP_H3, s.buf0 + MLP_COL_M1, MLP_LD0, s.wst);
      mlp_layer_nt_acc<MLP_H2>(acc, p + net.off[8], MLP_H3, MLP_H3, s.buf0 + MLP_COL_V1, MLP_LD0, s.wst);
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const int rr = rg * 8 + r, n = cg + 32 * j;
          float* y = s.buf1 + rr * s.ld1 + n;
          const float dz = acc[r][j] * elu_grad_from_out(*y);
          *y = dz;
          if (rr < nrows) w.dzh2[(size_t)(row0 + rr) * MLP_H2 + n] = dz;
        }
    }
    __syncthreads();
    {  // dZh1 = (dZh2 * W2) .* ELU'(h1); h1 was overwritten in smem -> re-read the saved copy (L2-resident)
      float acc[8][8];
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][j] = 0.0f;
      mlp_layer_nt_acc<MLP_H1>(acc, p + net.off[2], MLP_H2, MLP_H2, s.buf1, s.ld1, s.wst);
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const int rr = rg * 8 + r, n = cg + 32 * j;
          if (rr < nrows) {
            const size_t gi = (size_t)(row0 + rr) * MLP_H1 + n;
            w.dzh1[gi] = acc[r][j] * elu_grad_from_out(w.h1[gi]);
          }
        }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------ weight gradients
struct DwLayer {
  const float* dz;   // [rows][N]
  const float* x;    // [rows][ldx] (first K columns used)
  int N, K, ldx;
  int ldo;           // row stride of the output weight matrix
  int w_off, b_off;  // offsets in the flat gradient
  int tiles_n, tiles_k, tile0;
};
struct DwArgs {
  DwLayer layer[6];
  int total_tiles;
  long long rows;
  int rows_per_split;
  int net_total;
};

constexpr int DW_T = 64;    // output tile (n and k)
constexpr int DW_RC = 16;   // rows per smem stage

__global__ void __launch_bounds__(256)
ppo_dw_kernel(const __grid_constant__ DwArgs a, float* __restrict__ dw_part) {
  __shared__ __align__(16) float zs[DW_RC][DW_T];
  __shared__ __align__(16) float xs[DW_RC][DW_T];
  const int tile = blockIdx.x, split = blockIdx.y;
  int li = 0;
#pragma unroll
  for (int i = 1; i < 6; ++i)
    if (tile >= a.layer[i].tile0) li = i;
  const DwLayer L = a.layer[li];
  const int t = tile - L.tile0;
  const int tn = t / L.tiles_k, tk = t - tn * L.tiles_k;
  const int n0 = tn * DW_T, k0 = tk * DW_T;
  const long long r_begin = (long long)split * a.rows_per_split;
  const long long r_end = min(a.rows, r_begin + a.rows_per_split);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;   // tx -> k fragment, ty -> n fragment
  float acc[4][4];
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (long long r0 = r_begin; r0 < r_end; r0 += DW_RC) {
    // stage 16 rows x 64 columns of dZ and X (zero outside the matrices)
    for (int e = threadIdx.x; e < DW_RC * DW_T; e += 256) {
      const int rr = e >> 6, c = e & 63;
      const long long r = r0 + rr;
      const bool rok = r < r_end;
      zs[rr][c] = (rok && n0 + c < L.N) ? __ldg(L.dz + (size_t)r * L.N + n0 + c) : 0.0f;
      xs[rr][c] = (rok && k0 + c < L.K) ? __ldg(L.x + (size_t)r * L.ldx + k0 + c) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < DW_RC; ++rr) {
      const float4 z4 = *reinterpret_cast<const float4*>(&zs[rr][ty * 4]);
      const float4 x4 = *reinterpret_cast<const float4*>(&xs[rr][tx * 4]);
      const float zz[4] = {z4.x, z4.y, z4.z, z4.w};
      const float xx[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        bsum[i] += zz[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(zz[i], xx[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
  float* out = dw_part + (size_t)split * a.net_total;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int n = n0 + ty * 4 + i;
    if (n >= L.N) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + tx * 4 + j;
      if (k < L.ldo) out[L.w_off + (size_t)n * L.ldo + k] = acc[i][j];   // k in [K, ldo) holds exact zeros
    }
    if (tk == 0 && tx == 0) out[L.b_off + n] = bsum[i];
  }
}

template <bool PEERS>
__global__ void __launch_bounds__(256)
ppo_reduce_kernel(const float* __restrict__ dw_part, int splits, int total, float* __restrict__ grad,
                  const float* __restrict__ loss_part, long long ntiles, float inv_rows, float* __restrict__ loss,
                  const __grid_constant__ PeerView pv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t epoch = 0;
  if (PEERS) epoch = *(volatile const uint32_t*)pv.epoch;
  if (i < total) {
    float g = 0.0f;
    for (int s2 = 0; s2 < splits; ++s2) g += dw_part[(size_t)s2 * total + i];
    grad[i] = g;
    if (PEERS) peer_push(pv, epoch & 1u, i, g);
  }
  if (PEERS) peer_push_fence(pv);
  if (blockIdx.x == 0 && threadIdx.x < 32) {   // loss = mean(-surr) + smooth_l1 mean (ppo.py:194, 197)
    float x = 0.0f, y = 0.0f;
    for (long long tt = threadIdx.x; tt < ntiles; tt += 32) {
      x += loss_part[tt * 2 + 0];
      y += loss_part[tt * 2 + 1];
    }
    x = warp_sum(x);
    y = warp_sum(y);
    if (threadIdx.x == 0) *loss = -x * inv_rows + y * inv_rows;
  }
}

// ------------------------------------------------------------------------------------------ clip + Adam
// With PEERS the kernel first completes the data-parallel sum: it waits for every rank's push of this epoch, adds
// the `world` slots in rank order and stores the result to `grad` (what ppo_adam_kernel reads).
template <bool PEERS>
__global__ void __launch_bounds__(256)
ppo_norm_kernel(float* __restrict__ grad, int total, float scale, float* __restrict__ norm_part,
                long long* __restrict__ step, const __grid_constant__ PeerView pv) {
  __shared__ float red[8];
  uint32_t parity = 0;
  if (PEERS) {
    const uint32_t epoch = *(volatile const uint32_t*)pv.epoch;   // exchanges completed before this one
    if (blockIdx.x == 0) {   // the pushing kernel before this one is complete: its stores have reached every rank
      if (threadIdx.x == 0) pv.stamp[3] = global_ns();
      peer_raise_flags(pv, epoch + 1u);
    }
    peer_wait_all(pv, epoch + 1u);
    if (blockIdx.x == 0 && threadIdx.x == 0) pv.stamp[4] = global_ns();
    parity = epoch & 1u;
  }
  float acc = 0.0f;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    float g;
    if (PEERS) {
      g = peer_sum(pv, parity, i);
      grad[i] = g;
    } else {
      g = grad[i];
    }
    g *= scale;
    acc = fmaf(g, g, acc);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
    for (int i = 0; i < 8; ++i) t += red[i];
    norm_part[blockIdx.x] = t;
    if (blockIdx.x == 0) *step += 1;   // Adam's step counter; read by ppo_adam_kernel after this launch
    if (PEERS && blockIdx.x == 0) pv.stamp[5] = global_ns();
  }
}

__global__ void __launch_bounds__(256)
ppo_adam_kernel(float* __restrict__ params, float* __restrict__ grad, float* __restrict__ exp_avg,
                float* __restrict__ exp_avg_sq, int total, const float* __restrict__ norm_part,
                const long long* __restrict__ step, const __grid_constant__ AdamArgs a,
                float* __restrict__ grad_norm_out, uint32_t* __restrict__ peer_epoch, int norm_parts,
                float* __restrict__ wpack, const int4* __restrict__ wpack_index) {
  __shared__ AdamCoef s_c;
  if (threadIdx.x < 32) {
    float t = 0.0f;   // fixed order: lane l adds partials l, l + 32, ...; then the butterfly
    for (int i = threadIdx.x; i < norm_parts; i += 32) t += norm_part[i];
    t = warp_sum(t);
    if (threadIdx.x == 0) {
      s_c = adam_coefficients(a, t, *step);
      if (blockIdx.x == 0 && grad_norm_out) *grad_norm_out = s_c.total_norm;
      if (blockIdx.x == 0 && peer_epoch) *peer_epoch += 1;   // this exchange is over (every norm CTA has read the old value)
    }
  }
  __syncthreads();
  const AdamCoef c = s_c;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256)
    adam_entry(i, a, c, params, grad, exp_avg, exp_avg_sq, wpack, wpack_index);
}

template <typename K>
static int prepare_smem(K kernel, size_t smem) {
  return ensure_smem(reinterpret_cast<const void*>(kernel), smem);
}

}  // namespace flyb

using namespace flyb;

extern "C" size_t flyb_ppo_workspace_bytes(const FlybNetLayout* net, int64_t rows) {
  if (!net || rows <= 0) return 0;
  return std::max(carve_ws(*net, rows, nullptr).bytes, tc_ppo_workspace_bytes(net, rows));
}

extern "C" int flyb_ppo_grad(const FlybNetLayout* net, const FlybPpoCfg* cfg, const float* params, const void* wpack,
                             const float* obs, const float* action, const float* old_logp, const float* target,
                             const float* advantage, const float* action_var, int64_t rows, float* grad, float* loss,
                             void* workspace, const FlybPeerGroup* peers, void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg && obs && action && old_logp && target && advantage && action_var && grad && loss && workspace,
               "flyb_ppo_grad: NULL argument");
  FLYB_REQUIRE(rows > 0, "flyb_ppo_grad: rows must be > 0");
  FLYB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "flyb_ppo_grad: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  if (peers && (rc = peer_validate(peers, net->total)) != FLYB_OK) return rc;
  FLYB_REQUIRE(!wpack || tc_supported(net), "flyb_ppo_grad: wpack given but the tensor-core path does not cover this shape");
  if (wpack && tc_ppo_supported(net)) {   // tensor-core path (ppo_kernels_tc.cu); the wide shape's step runs below
    return tc_ppo_grad(net, cfg, params, wpack, obs, action, old_logp, target, advantage, action_var, rows, grad, loss,
                       workspace, peers, st, (StepFuse*)nullptr);
  }
  const PpoWs w = carve_ws(*net, rows, workspace);

  GradArgs a{};
  a.obs = obs; a.action = action; a.old_logp = old_logp; a.target = target; a.advantage = advantage;
  a.action_var = action_var; a.rows = rows;
  a.clip_lo = (float)(1.0 - cfg->clip);
  a.clip_hi = (float)(1.0 + cfg->clip);
  a.inv_rows = 1.0f / (float)rows;

  const size_t smem = mlp_smem_bytes(net->k0_pad);
  const int grid = (int)std::min<long long>(w.ntiles, (long long)num_sms());
  if (net->num_act <= 32) {
    if ((rc = prepare_smem(ppo_fwd_bwd_kernel<32>, smem)) != FLYB_OK) return rc;
    ppo_fwd_bwd_kernel<32><<<grid, MLP_THREADS, smem, st>>>(*net, params, a, w);
  } else {
    if ((rc = prepare_smem(ppo_fwd_bwd_kernel<64>, smem)) != FLYB_OK) return rc;
    ppo_fwd_bwd_kernel<64><<<grid, MLP_THREADS, smem, st>>>(*net, params, a, w);
  }
  FLYB_LAUNCH_CHECK("ppo_fwd_bwd_kernel");

  DwArgs d{};
  const int A = net->num_act, O = net->num_obs;
  const float* dzs[6] = {w.dzh1, w.dzh2, w.dzm1, w.dzmu, w.dzv1, w.dzval};
  const float* xsrc[6] = {obs, w.h1, w.h2, w.m1, w.h2, w.v1};
  const int Ns[6] = {MLP_H1, MLP_H2, MLP_H3, A, MLP_H3, 1};
  const int Ks[6] = {O, MLP_H1, MLP_H2, MLP_H3, MLP_H2, MLP_H3};
  const int ldos[6] = {net->k0_pad, MLP_H1, MLP_H2, MLP_H3, MLP_H2, MLP_H3};
  const int woff[6] = {net->off[0], net->off[2], net->off[4], net->off[6], net->off[8], net->off[10]};
  const int boff[6] = {net->off[1], net->off[3], net->off[5], net->off[7], net->off[9], net->off[11]};
  int tile0 = 0;
  for (int i = 0; i < 6; ++i) {
    DwLayer& L = d.layer[i];
    L.dz = dzs[i]; L.x = xsrc[i]; L.N = Ns[i]; L.K = Ks[i]; L.ldx = Ks[i]; L.ldo = ldos[i];
    L.w_off = woff[i]; L.b_off = boff[i];
    L.tiles_n = (Ns[i] + DW_T - 1) / DW_T;
    L.tiles_k = (ldos[i] + DW_T - 1) / DW_T;
    L.tile0 = tile0;
    tile0 += L.tiles_n * L.tiles_k;
  }
  d.total_tiles = tile0;
  d.rows = rows;
  d.rows_per_split = (int)(((rows + w.splits - 1) / w.splits + DW_RC - 1) / DW_RC * DW_RC);
  d.net_total = net->total;
  // segment padding (and nothing else) is left untouched by ppo_dw_kernel: clear the partials first
  rc = check_cuda(cudaMemsetAsync(w.dw_part, 0, (size_t)w.splits * net->total * sizeof(float), st), "cudaMemsetAsync");
  if (rc != FLYB_OK) return rc;
  ppo_dw_kernel<<<dim3(d.total_tiles, w.splits), 256, 0, st>>>(d, w.dw_part);
  FLYB_LAUNCH_CHECK("ppo_dw_kernel");
  if (peers)
    ppo_reduce_kernel<true><<<(net->total + 255) / 256, 256, 0, st>>>(w.dw_part, w.splits, net->total, grad, w.loss_part,
                                                                      w.ntiles, a.inv_rows, loss, make_peer_view(*peers));
  else
    ppo_reduce_kernel<false><<<(net->total + 255) / 256, 256, 0, st>>>(w.dw_part, w.splits, net->total, grad,
                                                                       w.loss_part, w.ntiles, a.inv_rows, loss, PeerView{});
  FLYB_LAUNCH_CHECK("ppo_reduce_kernel");
  return FLYB_OK;
}

// clip + Adam (+ in-place refresh of the tensor-core weight stream) on `grad`; norm partials come from whoever ran
// before: ppo_norm_kernel (flyb_ppo_apply) or the gradient reduction itself (flyb_ppo_step)
static int launch_adam(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, float* grad, float* exp_avg,
                       float* exp_avg_sq, const long long* step, float grad_scale, float* grad_norm_out,
                       const float* norm_part, int norm_parts, void* wpack, const FlybPeerGroup* peers,
                       cudaStream_t st) {
  if (wpack)
    FLYB_REQUIRE(tc_supported(net), "wpack given but the tensor-core path does not cover this shape");
  AdamArgs a{cfg->lr, cfg->beta1, cfg->beta2, cfg->eps, cfg->max_grad_norm, grad_scale};
  ppo_adam_kernel<<<(net->total + 255) / 256, 256, 0, st>>>(
      params, grad, exp_avg, exp_avg_sq, net->total, norm_part, step, a, grad_norm_out,
      peers ? make_peer_view(*peers).epoch : nullptr, norm_parts, (float*)wpack,
      wpack ? (const int4*)tc_wpack_index(net, wpack) : nullptr);
  FLYB_LAUNCH_CHECK("ppo_adam_kernel");
  return FLYB_OK;
}

extern "C" int flyb_ppo_apply(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, float* grad,
                              float* exp_avg, float* exp_avg_sq, int64_t* step, float grad_scale,
                              float* grad_norm_out, void* workspace, void* wpack, const FlybPeerGroup* peers,
                              void* stream) {
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg && grad && exp_avg && exp_avg_sq && step && workspace, "flyb_ppo_apply: NULL argument");
  FLYB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "flyb_ppo_apply: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  float* norm_part = (float*)workspace;   // first slot of the PPO workspace (see carve_ws)
  if (peers) {
    if ((rc = peer_validate(peers, net->total)) != FLYB_OK) return rc;
    ppo_norm_kernel<true><<<NORM_BLOCKS, 256, 0, st>>>(grad, net->total, grad_scale, norm_part, (long long*)step,
                                                       make_peer_view(*peers));
  } else {
    ppo_norm_kernel<false><<<NORM_BLOCKS, 256, 0, st>>>(grad, net->total, grad_scale, norm_part, (long long*)step,
                                                        PeerView{});
  }
  FLYB_LAUNCH_CHECK("ppo_norm_kernel");
  return launch_adam(net, cfg, params, grad, exp_avg, exp_avg_sq, (const long long*)step, grad_scale, grad_norm_out,
                     norm_part, NORM_BLOCKS, wpack, peers, st);
}

extern "C" int flyb_ppo_step_dp(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, void* wpack,
                                const float* obs, const float* action, const float* old_logp, const float* target,
                                const float* advantage, const float* action_var, int64_t rows, float* grad, float* loss,
                                float* exp_avg, float* exp_avg_sq, int64_t* step, float grad_scale, float* grad_norm_out,
                                void* workspace, const FlybPeerGroup* peers, void* stream) {
  FLYB_REQUIRE(peers, "flyb_ppo_step_dp: NULL peer group");
  const bool tc = wpack && tc_ppo_supported(net);
  // CUDA-core gradient, or a GPU too small to hold the fused exchange's reducing grid all at once (its CTAs wait for
  // the peers' CTAs of the same index: not resident = possible dead-lock): the two halves as they are (split exchange)
  if (!tc || !tc_ppo_exchange_resident(net)) {
    int rc0 = flyb_ppo_grad(net, cfg, params, tc ? wpack : nullptr, obs, action, old_logp, target, advantage, action_var,
                            rows, grad, loss, workspace, peers, stream);
    if (rc0 != FLYB_OK) return rc0;
    return flyb_ppo_apply(net, cfg, params, grad, exp_avg, exp_avg_sq, step, grad_scale, grad_norm_out, workspace, wpack,
                          peers, stream);
  }
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg && obs && action && old_logp && target && advantage && action_var && grad && loss && exp_avg &&
                   exp_avg_sq && step && workspace,
               "flyb_ppo_step_dp: NULL argument");
  FLYB_REQUIRE(rows > 0, "flyb_ppo_step_dp: rows must be > 0");
  FLYB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "flyb_ppo_step_dp: workspace must be 256-byte aligned");
  if ((rc = peer_validate(peers, net->total)) != FLYB_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  StepFuse fuse{};
  fuse.step = (long long*)step;
  const char* tail_env = getenv("FLYB_DP_TAIL");   // read per call: a test flips it
  fuse.tail = (tail_env && tail_env[0] == '1') ? 1 : 0;
  fuse.exchange = 1;
  fuse.params = params; fuse.exp_avg = exp_avg; fuse.exp_avg_sq = exp_avg_sq; fuse.grad_norm_out = grad_norm_out;
  fuse.lr = cfg->lr; fuse.beta1 = cfg->beta1; fuse.beta2 = cfg->beta2; fuse.eps = cfg->eps; fuse.max_norm = cfg->max_grad_norm;
  fuse.grad_scale = grad_scale;
  rc = tc_ppo_grad(net, cfg, params, wpack, obs, action, old_logp, target, advantage, action_var, rows, grad, loss,
                   workspace, peers, st, &fuse);
  if (rc != FLYB_OK) return rc;
  if (fuse.done == 1 || fuse.done == 3) return FLYB_OK;   // cooperative tail, or clip + Adam inside the reduction
  // fused exchange: the reduction left the sum over ranks in `grad` and its squared-norm partials; Adam closes the epoch
  return launch_adam(net, cfg, params, grad, exp_avg, exp_avg_sq, (const long long*)step, grad_scale, grad_norm_out,
                     fuse.sq_part, fuse.sq_parts, wpack, peers, st);
}

extern "C" int flyb_ppo_step(const FlybNetLayout* net, const FlybPpoCfg* cfg, float* params, void* wpack,
                             const float* obs, const float* action, const float* old_logp, const float* target,
                             const float* advantage, const float* action_var, int64_t rows, float* grad, float* loss,
                             float* exp_avg, float* exp_avg_sq, int64_t* step, float* grad_norm_out, void* workspace,
                             void* stream) {
  if (!wpack || !tc_ppo_supported(net)) {   // CUDA-core gradient: the two halves as they are (Adam still keeps a
    // forward-only weight stream — the wide shape's — in step)
    int rc = flyb_ppo_grad(net, cfg, params, nullptr, obs, action, old_logp, target, advantage, action_var, rows, grad,
                           loss, workspace, nullptr, stream);
    if (rc != FLYB_OK) return rc;
    return flyb_ppo_apply(net, cfg, params, grad, exp_avg, exp_avg_sq, step, 1.0f, grad_norm_out, workspace, wpack,
                          nullptr, stream);
  }
  int rc = validate_net(net, params);
  if (rc != FLYB_OK) return rc;
  FLYB_REQUIRE(cfg && obs && action && old_logp && target && advantage && action_var && grad && loss && exp_avg &&
                   exp_avg_sq && step && workspace,
               "flyb_ppo_step: NULL argument");
  FLYB_REQUIRE(rows > 0, "flyb_ppo_step: rows must be > 0");
  FLYB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "flyb_ppo_step: workspace must be 256-byte aligned");
  FLYB_REQUIRE(tc_ppo_supported(net), "flyb_ppo_step: wpack given but the tensor-core path does not cover this shape");
  cudaStream_t st = (cudaStream_t)stream;
  StepFuse fuse{};
  fuse.step = (long long*)step;
  // measured (profiles/r02e): the cooperative two-launch form is 3 us SLOWER per optimiser step than the four plain
  // launches (139.5 vs 136.2 us at 32768 samples) — two grid barriers over 148 heavy CTAs and a launch that cannot
  // overlap its predecessor's drain cost more than the two launch gaps they replace.  Opt-in: FLYB_PPO_TAIL=1.
  const char* tail_env = getenv("FLYB_PPO_TAIL");   // read per call: a test flips it
  const bool use_tail = tail_env && tail_env[0] == '1';
  fuse.tail = use_tail ? 1 : 0;
  fuse.params = params; fuse.exp_avg = exp_avg; fuse.exp_avg_sq = exp_avg_sq; fuse.grad_norm_out = grad_norm_out;
  fuse.lr = cfg->lr; fuse.beta1 = cfg->beta1; fuse.beta2 = cfg->beta2; fuse.eps = cfg->eps; fuse.max_norm = cfg->max_grad_norm;
  rc = tc_ppo_grad(net, cfg, params, wpack, obs, action, old_logp, target, advantage, action_var, rows, grad, loss,
                   workspace, nullptr, st, &fuse);
  if (rc != FLYB_OK) return rc;
  if (fuse.done) return FLYB_OK;   // cooperative tail (2 launches) or clip + Adam inside the reduction (3 launches)
  return launch_adam(net, cfg, params, grad, exp_avg, exp_avg_sq, (const long long*)step, 1.0f, grad_norm_out,
                     fuse.sq_part, fuse.sq_parts, wpack, nullptr, st);
}
