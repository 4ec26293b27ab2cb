// Kernel (c): rollout-buffer GAE / return computation as a reverse time scan, one thread per env.
//
// Replaces reference PPO.make_data (ppo.py:160-169): T sequential (N,1) launches become one; the two T*N-row
// value forwards collapse into one (T+1)*N-row forward because all_next_obs[t] == all_obs[t+1] (ppo.py:210,228).
//   target_t = r_t + gamma * v_{t+1} * done          done = 1 - reset_buf of the LAST step (ppo.py:230 quirk)
//   delta_t  = target_t - v_t
//   A_t      = gamma*lambda * A_{t+1} + delta_t      A_T = 0, no done mask inside the recursion
// Time-major (T,N) layout: at each t a warp reads 32 consecutive envs -> fully coalesced 128-byte lines.
// HBM traffic is 12 B read + 8 B written per (t, env) sample; v_{t+1} is kept in a register so each value is
// read once.  The t-loop is software-pipelined UNROLL deep so that every thread keeps 2*UNROLL loads in flight.
// When N is a multiple of 4 (and the buffers are 16-byte aligned) a thread owns FOUR adjacent envs and every load /
// store is a 128-bit vector access (gae_scan4_kernel): a quarter of the memory instructions for the same bytes; the
// per-env arithmetic, and so every result bit, is the same as in the scalar kernel.
#include "flyb_common.cuh"

namespace flyb {

constexpr int GAE_UNROLL = 8;

__global__ void __launch_bounds__(128)
gae_scan_kernel(const float* __restrict__ reward, const float* __restrict__ values,
                const long long* __restrict__ reset_last, int T, int N, float gamma, float gl,
                float* __restrict__ target, float* __restrict__ advantage) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= N) return;
  const float done = (float)(1 - reset_last[env]);
  float v_next = __ldg(values + (size_t)T * N + env);
  float adv = 0.0f;
  int t = T - 1;
  for (; t >= GAE_UNROLL - 1; t -= GAE_UNROLL) {
    float r[GAE_UNROLL], v[GAE_UNROLL];
#pragma unroll
    for (int u = 0; u < GAE_UNROLL; ++u) {
      r[u] = __ldcs(reward + (size_t)(t - u) * N + env);
      v[u] = __ldcs(values + (size_t)(t - u) * N + env);
    }
#pragma unroll
    for (int u = 0; u < GAE_UNROLL; ++u) {
      const float tg = __fadd_rn(r[u], __fmul_rn(__fmul_rn(gamma, v_next), done));
      const float delta = __fsub_rn(tg, v[u]);
      adv = __fadd_rn(__fmul_rn(gl, adv), delta);
      __stcs(target + (size_t)(t - u) * N + env, tg);
      __stcs(advantage + (size_t)(t - u) * N + env, adv);
      v_next = v[u];
    }
  }
  for (; t >= 0; --t) {
    const float r = __ldcs(reward + (size_t)t * N + env);
    const float v = __ldcs(values + (size_t)t * N + env);
    const float tg = __fadd_rn(r, __fmul_rn(__fmul_rn(gamma, v_next), done));
    const float delta = __fsub_rn(tg, v);
    adv = __fadd_rn(__fmul_rn(gl, adv), delta);
    __stcs(target + (size_t)t * N + env, tg);
    __stcs(advantage + (size_t)t * N + env, adv);
    v_next = v;
  }
}

FLYB_D float4 ldcs4(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }

__global__ void __launch_bounds__(64)
gae_scan4_kernel(const float* __restrict__ reward, const float* __restrict__ values,
                 const long long* __restrict__ reset_last, int T, int N, float gamma, float gl,
                 float* __restrict__ target, float* __restrict__ advantage) {
  constexpr int U = 4;   // 4 rows x (reward, value) x 16 bytes = 8 vector loads in flight per thread
  const int env = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (env >= N) return;
  float done[4], v_next[4], adv[4] = {0.f, 0.f, 0.f, 0.f};
  {
    const longlong2 r01 = *reinterpret_cast<const longlong2*>(reset_last + env);
    const longlong2 r23 = *reinterpret_cast<const longlong2*>(reset_last + env + 2);
    done[0] = (float)(1 - r01.x); done[1] = (float)(1 - r01.y); done[2] = (float)(1 - r23.x); done[3] = (float)(1 - r23.y);
    const float4 vn = ldcs4(values + (size_t)T * N + env);
    v_next[0] = vn.x; v_next[1] = vn.y; v_next[2] = vn.z; v_next[3] = vn.w;
  }
  auto row = [&](size_t at, const float4& r4, const float4& v4) {
    const float r[4] = {r4.x, r4.y, r4.z, r4.w}, v[4] = {v4.x, v4.y, v4.z, v4.w};
    float tg[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      tg[j] = __fadd_rn(r[j], __fmul_rn(__fmul_rn(gamma, v_next[j]), done[j]));
      const float delta = __fsub_rn(tg[j], v[j]);
      adv[j] = __fadd_rn(__fmul_rn(gl, adv[j]), delta);
      v_next[j] = v[j];
    }
    __stcs(reinterpret_cast<float4*>(target + at), make_float4(tg[0], tg[1], tg[2], tg[3]));
    __stcs(reinterpret_cast<float4*>(advantage + at), make_float4(adv[0], adv[1], adv[2], adv[3]));
  };
  int t = T - 1;
  for (; t >= U - 1; t -= U) {
    float4 r[U], v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      r[u] = ldcs4(reward + (size_t)(t - u) * N + env);
      v[u] = ldcs4(values + (size_t)(t - u) * N + env);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) row((size_t)(t - u) * N + env, r[u], v[u]);
  }
  for (; t >= 0; --t) row((size_t)t * N + env, ldcs4(reward + (size_t)t * N + env), ldcs4(values + (size_t)t * N + env));
}

}  // namespace flyb

using namespace flyb;

extern "C" int flyb_gae(const float* reward, const float* values, const int64_t* reset_last, int32_t horizon,
                        int32_t num_envs, float gamma, float lmbda, float* target, float* advantage, void* stream) {
  FLYB_REQUIRE(reward && values && reset_last && target && advantage, "flyb_gae: NULL buffer");
  FLYB_REQUIRE(horizon > 0 && num_envs > 0, "flyb_gae: horizon and num_envs must be > 0");
  // gamma * lmbda is formed in double then rounded, as Python does for `self.gamma * self.lmbda` (ppo.py:167)
  const float gl = (float)((double)gamma * (double)lmbda);
  // measured at configs[1] (16384 envs x 32): 4.3 us scalar vs 8.2 us vectorised — with a quarter of the threads the
  // scan is latency-bound; the vector kernel takes over once N / 4 threads still fill the 148 SMs
  const bool vec = (num_envs % 4 == 0) && num_envs >= 4 * 148 * 256 && aligned16(reward) && aligned16(values) &&
                   aligned16(target) && aligned16(advantage) && aligned16(reset_last);
  if (vec) {
    gae_scan4_kernel<<<(num_envs / 4 + 63) / 64, 64, 0, (cudaStream_t)stream>>>(
        reward, values, (const long long*)reset_last, horizon, num_envs, gamma, gl, target, advantage);
  } else {
    gae_scan_kernel<<<(num_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        reward, values, (const long long*)reset_last, horizon, num_envs, gamma, gl, target, advantage);
  }
  FLYB_LAUNCH_CHECK("gae_scan_kernel");
  return FLYB_OK;
}
