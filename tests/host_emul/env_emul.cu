// TEST-ONLY host build of the per-env device functions (fly_bproject_b200/csrc/env_math.cuh).
//
// There is no GPU in the build container; this shim lets the CPU test-suite run the very same
// __host__ __device__ arithmetic the CUDA kernel executes against the golden vectors before any GPU time is
// spent.  It is NOT part of libflyb200.so, is never imported by the package, and is not a fallback: the product
// path has no CPU implementation.
#include "../../fly_bproject_b200/csrc/env_math.cuh"

using namespace flyb;

extern "C" void emul_env_post_physics(const FlybEnvCfg* cfg, const float* root, const float* dof_state,
                                      const float* force, const float* actions_scaled, float* potentials,
                                      float* prev_potentials, long long* progress, long long* reset, float* obs,
                                      float* reward, float* up_vec, float* heading_vec) {
  const FlybEnvCfg& c = *cfg;
  const int N = c.num_envs, D = c.num_dof, A = c.num_act, B = c.num_bodies, O = 19 + 3 * A;
  for (int env = 0; env < N; ++env) {
    float* o = obs + (size_t)env * O;
    for (int k = 0; k < A; ++k) {
      const int d = c.action_idx[k];
      o[12 + k] = unscale(dof_state[((size_t)env * D + d) * 2], c.lower[d], c.upper[d]);
      o[12 + A + k] = dof_state[((size_t)env * D + d) * 2 + 1] * c.dof_vel_scale;
      o[12 + 2 * A + k] = actions_scaled[(size_t)env * A + k];
    }
    float f[FORCE_VALS];
    for (int r = 0; r < FORCE_ROWS; ++r) {
      const int body = (r < FLYB_NUM_ABDOMEN) ? c.abdomen_local[r] : c.legtip_local[r - FLYB_NUM_ABDOMEN];
      for (int comp = 0; comp < 3; ++comp) f[r * 3 + comp] = force[((size_t)env * B + body) * 3 + comp];
    }
    const EnvStepOut out = env_post_one(c, root + (size_t)env * 13, f, o, potentials[env], progress[env], reset[env]);
    potentials[env] = out.potential;
    prev_potentials[env] = out.prev_potential;
    progress[env] = out.progress;
    reset[env] = out.reset;
    reward[env] = out.reward;
    up_vec[env * 3 + 0] = out.up_vec.x; up_vec[env * 3 + 1] = out.up_vec.y; up_vec[env * 3 + 2] = out.up_vec.z;
    heading_vec[env * 3 + 0] = out.heading_vec.x; heading_vec[env * 3 + 1] = out.heading_vec.y;
    heading_vec[env * 3 + 2] = out.heading_vec.z;
  }
}

extern "C" void emul_env_scale_actions(const FlybEnvCfg* cfg, const float* actions, float* actions_scaled,
                                       float* dof_targets) {
  const FlybEnvCfg& c = *cfg;
  for (int env = 0; env < c.num_envs; ++env) {
    for (int d = 0; d < c.num_dof; ++d) dof_targets[(size_t)env * c.num_dof + d] = c.initial_pos[d];
    for (int k = 0; k < c.num_act; ++k) {
      const int d = c.action_idx[k];
      const float v = scale_action(actions[(size_t)env * c.num_act + k], c.lower[d], c.upper[d]);
      actions_scaled[(size_t)env * c.num_act + k] = v;
      dof_targets[(size_t)env * c.num_dof + d] = v;
    }
  }
}
