// Per-env arithmetic of the fly observation / reward / termination path.
//
// Follows, in the reference's operation order:
//   compute_fly_observations   reference fly.py:771-805
//   compute_fly_reward2        reference fly.py:685-768
// and the un-vendored Isaac Gym helpers they call (isaacgym.torch_utils / isaacgymenvs torch_jit_utils:
// quat_mul, quat_rotate, quat_rotate_inverse, normalize, get_euler_xyz, unscale, scale — formulas in
// SURVEY.md Appendix A; call sites fly.py:629, 789-795).
// Quaternions are (x, y, z, w).  Everything is fp32; no FMA contraction is relied on either way.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

struct Vec3 {
  float x, y, z;
};
struct Quat {
  float x, y, z, w;
};

FLYB_HD Quat quat_mul(const Quat& a, const Quat& b) {
  const float ww = (a.z + a.x) * (b.x + b.y);
  const float yy = (a.w - a.y) * (b.w + b.z);
  const float zz = (a.w + a.y) * (b.w - b.z);
  const float xx = ww + yy + zz;
  const float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
  Quat r;
  r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
  r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
  r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
  r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
  return r;
}

// quat_rotate (sign = +1) / quat_rotate_inverse (sign = -1): v*(2w^2-1) +/- cross(q.xyz, v)*w*2 + q.xyz*dot(q.xyz,v)*2
FLYB_HD Vec3 quat_rotate_signed(const Quat& q, const Vec3& v, float sign) {
  const float s = 2.0f * (q.w * q.w) - 1.0f;
  const float cx = q.y * v.z - q.z * v.y;
  const float cy = q.z * v.x - q.x * v.z;
  const float cz = q.x * v.y - q.y * v.x;
  const float d = q.x * v.x + q.y * v.y + q.z * v.z;
  Vec3 r;
  r.x = v.x * s + sign * (cx * q.w * 2.0f) + q.x * d * 2.0f;
  r.y = v.y * s + sign * (cy * q.w * 2.0f) + q.y * d * 2.0f;
  r.z = v.z * s + sign * (cz * q.w * 2.0f) + q.z * d * 2.0f;
  return r;
}

// torch.remainder(x, 2*pi) for fp32: fmod, then shift negatives up by the (positive) divisor
FLYB_HD float wrap_2pi(float x) {
  const float two_pi = 6.283185307179586f;
  float m = fmodf(x, two_pi);
  if (m != 0.0f && m < 0.0f) m += two_pi;
  return m;
}

struct EnvObsHead {
  float height;
  Vec3 vel_loc, angvel_loc;
  float yaw, roll, angle_to_target, up_proj, heading_proj, pitch;
  Vec3 up_vec, heading_vec;
  float potential;
};

// root = [pos3, quat4 (xyzw), linvel3, angvel3]  (reference fly.py:95-100).
// The head of the observation is split in three independent pieces so that the kernel can give each to a different
// thread of an env; env_obs_head() chains them for the one-thread-per-env users (host emulation, tests).
FLYB_HD Quat env_torso_quat(const float* root, const FlybEnvCfg& c) {
  const Quat q = {root[3], root[4], root[5], root[6]};
  const Quat isr = {c.inv_start_rot[0], c.inv_start_rot[1], c.inv_start_rot[2], c.inv_start_rot[3]};
  return quat_mul(q, isr);
}
// |v|^2 the way torch's CPU norm forms it (fly.py:781): every product and every sum rounded on its own.  nvcc would
// contract x*x + y*y into an FMA, which moves the potential (O(6e4), ulp 2^-8) by an ulp in ~10 % of the envs — and the
// walking total holds 2 * (potential - previous potential).  With this form the golden potentials are reproduced bit
// for bit (sqrt and the division by dt are IEEE-exact on both sides).  Host builds use -ffp-contract=off.
FLYB_HD float sumsq3_rn(float x, float y, float z) {
#ifdef __CUDA_ARCH__
  return __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z));
#else
  const float a = x * x, b = y * y, c = z * z;
  return (a + b) + c;
#endif
}
// potential, compute_heading_and_up (fly.py:778-783, 789-790)
FLYB_HD void env_head_heading(const float* root, const FlybEnvCfg& c, const Quat& tq, EnvObsHead& o) {
  const Vec3 pos = {root[0], root[1], root[2]};
  Vec3 tt = {c.target[0] - pos.x, c.target[1] - pos.y, 0.0f};
  const float nrm = sqrtf(sumsq3_rn(tt.x, tt.y, tt.z));
  o.potential = -nrm / c.dt;
  const float nc = fmaxf(nrm, 1e-9f);
  const Vec3 dir = {tt.x / nc, tt.y / nc, tt.z / nc};
  const Vec3 b1 = {c.basis1[0], c.basis1[1], c.basis1[2]};
  const Vec3 b0 = {c.basis0[0], c.basis0[1], c.basis0[2]};
  o.up_vec = quat_rotate_signed(tq, b1, 1.0f);
  o.heading_vec = quat_rotate_signed(tq, b0, 1.0f);
  o.up_proj = o.up_vec.z;  // up_axis_idx = 2 (fly.py:17)
  o.heading_proj = o.heading_vec.x * dir.x + o.heading_vec.y * dir.y + o.heading_vec.z * dir.z;
}
// compute_rot, velocities (fly.py:792-793)
FLYB_HD void env_head_vel(const float* root, const Quat& tq, EnvObsHead& o) {
  const Vec3 lin = {root[7], root[8], root[9]};
  const Vec3 ang = {root[10], root[11], root[12]};
  o.vel_loc = quat_rotate_signed(tq, lin, -1.0f);
  o.angvel_loc = quat_rotate_signed(tq, ang, -1.0f);
  o.height = root[2];
}
// compute_rot, Euler angles and angle to target (fly.py:792-793)
FLYB_HD void env_head_euler(const float* root, const FlybEnvCfg& c, const Quat& tq, EnvObsHead& o) {
  const Vec3 pos = {root[0], root[1], root[2]};
  const float sinr_cosp = 2.0f * (tq.w * tq.x + tq.y * tq.z);
  const float cosr_cosp = tq.w * tq.w - tq.x * tq.x - tq.y * tq.y + tq.z * tq.z;
  const float roll = atan2f(sinr_cosp, cosr_cosp);
  const float sinp = 2.0f * (tq.w * tq.y - tq.z * tq.x);
  const float half_pi = 1.5707963267948966f;
  const float sgn = (sinp > 0.0f) ? 1.0f : ((sinp < 0.0f) ? -1.0f : 0.0f);
  const float pitch = (fabsf(sinp) >= 1.0f) ? half_pi * sgn : asinf(sinp);
  const float siny_cosp = 2.0f * (tq.w * tq.z + tq.x * tq.y);
  const float cosy_cosp = tq.w * tq.w + tq.x * tq.x - tq.y * tq.y - tq.z * tq.z;
  const float yaw = atan2f(siny_cosp, cosy_cosp);
  o.roll = wrap_2pi(roll);
  o.pitch = wrap_2pi(pitch);
  o.yaw = wrap_2pi(yaw);
  // walk_target_angle uses the z component, as upstream
  const float walk_target_angle = atan2f(c.target[2] - pos.z, c.target[0] - pos.x);
  o.angle_to_target = walk_target_angle - o.yaw;
}
FLYB_HD EnvObsHead env_obs_head(const float* root, const FlybEnvCfg& c) {
  EnvObsHead o;
  const Quat tq = env_torso_quat(root, c);
  env_head_heading(root, c, tq, o);
  env_head_vel(root, tq, o);
  env_head_euler(root, c, tq, o);
  return o;
}

FLYB_HD float unscale(float x, float lo, float hi) { return (2.0f * x - hi - lo) / (hi - lo); }
FLYB_HD float scale_action(float a, float lo, float hi) { return 0.5f * (a + 1.0f) * (hi - lo) + lo; }

struct EnvRewardIn {
  float height, heading_proj;
  float qz, qw;               // root orientation components (fly.py:725)
  float potential, prev_potential;
  float actions_sq_sum;       // sum(actions^2)                           (fly.py:731)
  float electricity;          // sum(|actions - obs[12+2A:12+3A]|)  == 0  (fly.py:733)
  int at_limit;               // #(cmd > 0.9*upper) + #(cmd < 0.9*lower)  (fly.py:736-737)
  int legs_touching;          // #(fx+fy+fz > 0) over the six leg tips    (fly.py:744)
  bool abdomen_hit;           // sum over 5 abdomen bodies of (fx+fy+fz) > 0 (fly.py:756)
  long long progress;         // after += 1 (fly.py:678)
  long long reset_prev;
};

FLYB_HD void env_reward_reset(const EnvRewardIn& in, const FlybEnvCfg& c, float* reward, long long* reset) {
  // fly.py:715-716
  const float heading_reward =
      (in.heading_proj > 0.8f) ? c.heading_weight : c.heading_weight * in.heading_proj / 0.8f;
  // fly.py:719-721
  float up_reward = 0.0f;
  if (in.height > 1.4f) up_reward = up_reward + c.up_weight;
  if (in.height < 2.1f) up_reward = up_reward - c.up_weight;
  // fly.py:725-728
  const float planar = in.qz * in.qz + in.qw * in.qw;
  const float orient_reward = (planar > 0.98f) ? c.up_weight : 0.0f;
  const float alive = 0.5f;                                 // fly.py:740
  const float legs = (float)in.legs_touching * 0.1f;        // fly.py:744
  const float limit_cost = (float)in.at_limit * c.joints_at_limit_cost_scale;
  float total;
  if (c.walking) {
    // fly.py:747-748 (commented-out walking total) + leg term (FinalReportTimo.pdf p.7)
    const float progress_reward = in.potential - in.prev_potential;
    total = progress_reward * 2.0f + alive + up_reward * orient_reward + heading_reward -
            c.actions_cost_scale * in.actions_sq_sum - c.energy_cost_scale * in.electricity - limit_cost + legs;
  } else {
    // fly.py:750
    total = alive + up_reward * orient_reward - c.energy_cost_scale * in.electricity - limit_cost + legs;
  }
  // fly.py:753-756
  const bool too_low = in.height < c.termination_height;
  const bool too_high = in.height > c.termination_height_up;
  const bool flipped = planar < 0.5f;
  if (too_low || too_high || flipped || in.abdomen_hit) total = c.death_cost;
  // fly.py:759-766
  const bool timeout = (float)in.progress >= c.max_episode_length - 1.0f;
  *reset = (too_low || too_high || timeout || flipped || in.abdomen_hit) ? 1 : in.reset_prev;
  *reward = total;
}


constexpr int FORCE_ROWS = FLYB_NUM_ABDOMEN + FLYB_NUM_LEGTIPS;  // 11 gathered contact-force rows per env
constexpr int FORCE_VALS = FORCE_ROWS * 3;                       // 33 (odd: conflict-free per-env reads)

struct EnvStepOut {
  float potential, prev_potential, reward;
  long long progress, reset;
  Vec3 up_vec, heading_vec;
};

// One env of get_obs + progress += 1 + get_reward (fly.py:676-681).  `o` is the env's observation row whose
// slots [12 : 12+3A] (scaled DOF pos, scaled DOF vel, commanded actions) are already filled; `f` holds the 11
// gathered force rows (5 abdomen, then 6 leg tips).  `o` may live in shared memory with any row stride.
FLYB_HD EnvStepOut env_post_one(const FlybEnvCfg& c, const float* r, const float* f, float* o, float potential_in,
                                long long progress_in, long long reset_in) {
  const int A = c.num_act;
  const EnvObsHead h = env_obs_head(r, c);
  o[0] = h.height;
  o[1] = h.vel_loc.x; o[2] = h.vel_loc.y; o[3] = h.vel_loc.z;
  o[4] = h.angvel_loc.x; o[5] = h.angvel_loc.y; o[6] = h.angvel_loc.z;
  o[7] = h.yaw; o[8] = h.roll; o[9] = h.angle_to_target; o[10] = h.up_proj; o[11] = h.heading_proj;
  o[12 + 3 * A] = h.pitch;

  float abdomen = 0.0f;
#pragma unroll
  for (int b = 0; b < FLYB_NUM_ABDOMEN; ++b) abdomen += (f[b * 3] + f[b * 3 + 1]) + f[b * 3 + 2];
  int legs = 0;
#pragma unroll
  for (int b = 0; b < FLYB_NUM_LEGTIPS; ++b) {
    const float* fr = f + (FLYB_NUM_ABDOMEN + b) * 3;
    const bool touch = ((fr[0] + fr[1]) + fr[2]) > 0.0f;
    o[13 + 3 * A + b] = touch ? 1.0f : 0.0f;  // fly.py:797, 803
    legs += touch ? 1 : 0;
  }

  EnvRewardIn in;
  in.actions_sq_sum = 0.0f;
  in.electricity = 0.0f;
  in.at_limit = 0;
  const float* cmd = o + 12 + 2 * A;
  for (int k = 0; k < A; ++k) {
    const float a = cmd[k];
    const int d = c.action_idx[k];
    in.actions_sq_sum += a * a;
    in.electricity += fabsf(a - a);  // fly.py:733: the obs slice IS `actions`, so this is identically 0
    in.at_limit += (a > c.upper[d] * 0.9f) ? 1 : 0;
    in.at_limit += (a < c.lower[d] * 0.9f) ? 1 : 0;
  }
  in.height = h.height;
  in.heading_proj = h.heading_proj;
  in.qz = r[5];
  in.qw = r[6];
  in.prev_potential = potential_in;  // fly.py:780
  in.potential = h.potential;
  in.legs_touching = legs;
  in.abdomen_hit = abdomen > 0.0f;
  long long prog = progress_in + 1;  // fly.py:678
  if (prog == 0) prog = 1;           // fly.py:415-416
  in.progress = prog;
  in.reset_prev = reset_in;
  EnvStepOut out;
  env_reward_reset(in, c, &out.reward, &out.reset);
  out.potential = h.potential;
  out.prev_potential = potential_in;
  out.progress = prog;
  out.up_vec = h.up_vec;
  out.heading_vec = h.heading_vec;
  return out;
}

}  // namespace flyb
