// Host-side dispatch between the two compiled shapes of the tensor-core kernels (mlp_tc.cuh): flyb::tcn (narrow:
// num_obs in (64, 80], num_act <= 32 — forward, act, fused env step and the PPO step) and flyb::tcw (wide: num_obs <= 160,
// num_act <= 48 — forward, act; the PPO step where the padded first-layer pitch is exactly 160, i.e. num_obs in (144, 160]).
#pragma once
#include "flyb_common.cuh"

namespace flyb {

#define FLYB_TC_DECLARE(NS)                                                                                              \
  namespace NS {                                                                                                         \
  bool tc_supported(const FlybNetLayout* net);                                                                           \
  bool tc_act_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg);                                                \
  size_t tc_wpack_bytes(const FlybNetLayout* net);                                                                       \
  const int* tc_wpack_index(const FlybNetLayout* net, const void* wpack);                                                \
  int tc_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, cudaStream_t st);                      \
  int tc_launch_forward(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs,             \
                        long long rows, float* mu_out, float* value_out, cudaStream_t st);                               \
  int tc_launch_act(const FlybNetLayout* net, const float* params, const void* wpack, const float* obs, long long rows,  \
                    const float* action_var, const float* noise, uint64_t seed, uint64_t offset,                         \
                    const uint64_t* offset_dev, float* action, float* logp, float* mu_out, float* obs_store,             \
                    const FlybEnvCfg* cfg, float* actions_scaled, float* dof_targets, int64_t* reset, int64_t* progress, \
                    float* potentials, float* prev_potentials, float* root, float* dof_state, cudaStream_t st);          \
  }
FLYB_TC_DECLARE(tcn)
FLYB_TC_DECLARE(tcw)
#undef FLYB_TC_DECLARE

struct StepFuse;
#define FLYB_TC_DECLARE_PPO(NS)                                                                                          \
  namespace NS {                                                                                                         \
  bool tc_ppo_supported(const FlybNetLayout* net);                                                                       \
  size_t tc_ppo_workspace_bytes(const FlybNetLayout* net, long long rows);                                               \
  bool tc_ppo_exchange_resident(const FlybNetLayout* net);                                                               \
  int tc_ppo_grad(const FlybNetLayout* net, const FlybPpoCfg* cfg, const float* params, const void* wpack,              \
                  const float* obs, const float* action, const float* old_logp, const float* target,                     \
                  const float* advantage, const float* action_var, long long rows, float* grad, float* loss,             \
                  void* workspace, const FlybPeerGroup* peers, cudaStream_t st, StepFuse* fuse);                         \
  }
FLYB_TC_DECLARE_PPO(tcn)
FLYB_TC_DECLARE_PPO(tcw)
#undef FLYB_TC_DECLARE_PPO

namespace tcn {
bool tc_step_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg);
int tc_launch_step(const FlybNetLayout* net, const float* params, const void* wpack, const FlybEnvCfg* cfg,
                   const FlybStepIO* io, cudaStream_t st);
}  // namespace tcn

inline bool tc_wide(const FlybNetLayout* net) { return !tcn::tc_supported(net) && tcw::tc_supported(net); }
// the forward / act kernels cover the shape (either compiled shape)
inline bool tc_supported(const FlybNetLayout* net) { return tcn::tc_supported(net) || tcw::tc_supported(net); }
// the tensor-core PPO step (chain + weight-gradient kernels) covers the shape
inline bool tc_ppo_supported(const FlybNetLayout* net) { return tcn::tc_ppo_supported(net) || tcw::tc_ppo_supported(net); }
inline size_t tc_ppo_workspace_bytes(const FlybNetLayout* net, long long rows) {
  return tcn::tc_ppo_supported(net) ? tcn::tc_ppo_workspace_bytes(net, rows) : tcw::tc_ppo_workspace_bytes(net, rows);
}
// the fused gradient exchange's reducing grid fits the GPU all at once (it must: see ppo_kernels_tc.cu)
inline bool tc_ppo_exchange_resident(const FlybNetLayout* net) {
  return tcn::tc_ppo_supported(net) ? tcn::tc_ppo_exchange_resident(net) : tcw::tc_ppo_exchange_resident(net);
}
template <class... Args>
inline int tc_ppo_grad(const FlybNetLayout* net, Args... args) {
  return tcn::tc_ppo_supported(net) ? tcn::tc_ppo_grad(net, args...) : tcw::tc_ppo_grad(net, args...);
}
inline bool tc_act_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg) {
  return tc_wide(net) ? tcw::tc_act_supported(net, cfg) : tcn::tc_act_supported(net, cfg);
}
inline bool tc_step_supported(const FlybNetLayout* net, const FlybEnvCfg* cfg) { return tcn::tc_step_supported(net, cfg); }
inline size_t tc_wpack_bytes(const FlybNetLayout* net) {
  return tcn::tc_supported(net) ? tcn::tc_wpack_bytes(net) : tcw::tc_wpack_bytes(net);
}
inline const int* tc_wpack_index(const FlybNetLayout* net, const void* wpack) {
  return tc_wide(net) ? tcw::tc_wpack_index(net, wpack) : tcn::tc_wpack_index(net, wpack);
}
inline int tc_pack_weights(const FlybNetLayout* net, const float* params, void* wpack, cudaStream_t st) {
  return tc_wide(net) ? tcw::tc_pack_weights(net, params, wpack, st) : tcn::tc_pack_weights(net, params, wpack, st);
}
template <class... Args>
inline int tc_launch_forward(const FlybNetLayout* net, Args... args) {
  return tc_wide(net) ? tcw::tc_launch_forward(net, args...) : tcn::tc_launch_forward(net, args...);
}
template <class... Args>
inline int tc_launch_act(const FlybNetLayout* net, Args... args) {
  return tc_wide(net) ? tcw::tc_launch_act(net, args...) : tcn::tc_launch_act(net, args...);
}
inline int tc_launch_step(const FlybNetLayout* net, const float* params, const void* wpack, const FlybEnvCfg* cfg,
                          const FlybStepIO* io, cudaStream_t st) {
  return tcn::tc_launch_step(net, params, wpack, cfg, io, st);
}

}  // namespace flyb
