"""CPU (torch fp32) restatement of the reference hot path — the travelling oracle.

TEST INFRASTRUCTURE (see oracle/__init__.py): checker for the CUDA path and the
timed CPU baseline of ``bench.py``; never on the product path.

Each function cites the reference lines it follows.  Arithmetic is kept in the
reference's operation order so that, on the same machine, results are
bit-identical to the executed reference (``tests/test_oracle_vs_reference.py``
asserts that where ``/root/reference`` exists; ``tests/golden`` freezes the
vectors for machines where it does not).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor

from . import isaac_math as im

# --------------------------------------------------------------------------- constants
# reference fly.py:16-51 and the literals inside compute_fly_reward2 (fly.py:716-744)
DT = 1.0 / 60.0
DOF_VEL_SCALE = 0.2
HEADING_WEIGHT = 0.5
UP_WEIGHT = 0.75
ACTIONS_COST_SCALE = 0.005
ENERGY_COST_SCALE = 0.005
JOINTS_AT_LIMIT_COST_SCALE = 0.1
DEATH_COST = -2.0
TERMINATION_HEIGHT = 1.1
TERMINATION_HEIGHT_UP = 6.0
MAX_EPISODE_LENGTH = 1500.0
STARTING_HEIGHT = 2.0
TARGET = (1000.0, 0.0, 0.0)

# reference ppo.py:113-119,152,237
GAMMA = 0.99
LMBDA = 0.95
CLIP = 0.2
LR = 1e-3
EPOCHS = 5
MINI_BATCH_SIZE = 40960
CHUNK_NUMBER = 16
VAR_START = 0.2
VAR_FLOOR = 0.01
VAR_DECAY = 0.00001


# --------------------------------------------------------------------------- env: observations
@torch.jit.script
def observations(root_states: Tensor, targets: Tensor, potentials: Tensor, inv_start_rot: Tensor,
                 dof_pos: Tensor, dof_vel: Tensor, lower: Tensor, upper: Tensor, dof_vel_scale: float,
                 actions: Tensor, dt: float, basis0: Tensor, basis1: Tensor, action_idx: Tensor,
                 force: Tensor, leg_tip_ids: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """reference fly.py:771-805 (compute_fly_observations).

    ``action_idx`` is (A,1) int64 like the reference's ``action_indexes_one``; ``leg_tip_ids`` is the
    flat (6N,) sim-domain id list.  Returns obs, potentials, prev_potentials, up_vec, heading_vec.
    """
    n = root_states.shape[0]
    pos = root_states[:, 0:3]
    quat = root_states[:, 3:7]
    lin = root_states[:, 7:10]
    ang = root_states[:, 10:13]

    to_target = targets - pos
    to_target[:, 2] = 0.0
    prev_potentials = potentials.clone()
    new_potentials = -torch.norm(to_target, p=2, dim=-1) / dt

    torso_quat, up_proj, heading_proj, up_vec, heading_vec = im.compute_heading_and_up(
        quat, inv_start_rot, to_target, basis0, basis1, 2)
    vel_loc, angvel_loc, roll, pitch, yaw, angle_to_target = im.compute_rot(
        torso_quat, lin, ang, targets, pos)

    pos_scaled = im.unscale(dof_pos, lower, upper)
    tip_sum = torch.sum(force[leg_tip_ids, :], dim=1).view(n, -1)
    touching = (tip_sum > 0).long()

    obs = torch.cat((pos[:, 2].view(-1, 1), vel_loc, angvel_loc, yaw.unsqueeze(-1), roll.unsqueeze(-1),
                     angle_to_target.unsqueeze(-1), up_proj.unsqueeze(-1), heading_proj.unsqueeze(-1),
                     pos_scaled[:, action_idx].squeeze(-1), dof_vel[:, action_idx].squeeze(-1) * dof_vel_scale,
                     actions, pitch.unsqueeze(-1), touching), dim=-1)
    return obs, new_potentials, prev_potentials, up_vec, heading_vec


# --------------------------------------------------------------------------- env: reward + reset mask
@torch.jit.script
def reward_reset(obs: Tensor, reset_buf: Tensor, progress_buf: Tensor, actions: Tensor, potentials: Tensor,
                 prev_potentials: Tensor, orientation: Tensor, force: Tensor, abdomen_ids: Tensor,
                 leg_tip_ids: Tensor, action_idx: Tensor, upper: Tensor, lower: Tensor, num_actions: int,
                 up_weight: float, heading_weight: float, actions_cost_scale: float, energy_cost_scale: float,
                 joints_at_limit_cost_scale: float, termination_height: float, termination_height_up: float,
                 death_cost: float, max_episode_length: float, walking: bool,
                 walking_leg_term: bool = True) -> Tuple[Tensor, Tensor]:
    """reference fly.py:685-768 (compute_fly_reward2).

    ``walking=False`` is the live standing total (fly.py:750).  ``walking=True`` is the commented-out
    walking total of fly.py:747-748 — pinned by executing the reference's function with those two lines switched back on
    (ref_harness.load_walking_reward, tests/test_oracle_vs_reference.py).  Inside that dead line the leg term is
    commented out once more (``#+ leg_ground_reward``): ``walking_leg_term=True`` (default; what the CUDA kernels
    compute) re-enables it too, an interpretation that follows BASELINE.json configs[2] / the report (p.7), fixture
    tests/golden/env_walking_d18_a18_n211.npz; ``walking_leg_term=False`` is the literal un-commenting, fixture
    env_walking_literal_d18_a18_n211.npz.
    """
    n = obs.shape[0]
    height = obs[:, 0]
    heading_proj = obs[:, 11]

    heading_w = torch.ones_like(heading_proj) * heading_weight
    heading_reward = torch.where(heading_proj > 0.8, heading_w, heading_weight * heading_proj / 0.8)

    up_reward = torch.zeros_like(heading_reward)
    up_reward = torch.where(height > 1.4, up_reward + up_weight, up_reward)
    up_reward = torch.where(height < 2.1, up_reward - up_weight, up_reward)

    planar = torch.square(orientation[:, 2]) + torch.square(orientation[:, 3])
    orient_reward = torch.zeros_like(up_reward)
    orient_reward = torch.where(planar > 0.98, orient_reward + up_weight, orient_reward)

    start = 12 + 2 * num_actions
    commanded = obs[:, start:(start + num_actions)]
    actions_cost = torch.sum(actions ** 2, dim=-1)
    electricity_cost = torch.sum(torch.abs(actions - commanded), dim=-1)

    hi = upper[action_idx].squeeze(-1).repeat((n, 1)) * 0.9
    lo = lower[action_idx].squeeze(-1).repeat((n, 1)) * 0.9
    at_limit = torch.sum(commanded > hi, dim=-1)
    at_limit += torch.sum(commanded < lo, dim=-1)

    alive_reward = torch.ones_like(potentials) * 0.5
    progress_reward = potentials - prev_potentials

    legs = torch.sum((torch.sum(force[leg_tip_ids, :], dim=1).view(n, -1) > 0).long(), dim=1) * 0.1

    if walking:
        total = progress_reward * 2 + alive_reward + up_reward * orient_reward + heading_reward - \
            actions_cost_scale * actions_cost - energy_cost_scale * electricity_cost - \
            at_limit * joints_at_limit_cost_scale
        if walking_leg_term:
            total = total + legs
    else:
        total = alive_reward + up_reward * orient_reward - energy_cost_scale * electricity_cost - \
            at_limit * joints_at_limit_cost_scale + legs

    abdomen_hit = torch.sum(torch.sum(force[abdomen_ids, :], dim=1).view(n, -1), dim=1) > 0
    dead = torch.ones_like(total) * death_cost
    total = torch.where(height < termination_height, dead, total)
    total = torch.where(height > termination_height_up, dead, total)
    total = torch.where(planar < 0.5, dead, total)
    total = torch.where(abdomen_hit, dead, total)

    one = torch.ones_like(reset_buf)
    reset = torch.where(height < termination_height, one, reset_buf)
    reset = torch.where(height > termination_height_up, one, reset)
    reset = torch.where(progress_buf >= max_episode_length - 1, one, reset)
    reset = torch.where(planar < 0.5, one, reset)
    reset = torch.where(abdomen_hit, one, reset)
    return total, reset


# --------------------------------------------------------------------------- env state container
class EnvTables:
    """The index tables and constants ``Fly.__init__`` derives (reference fly.py:248-348)."""

    def __init__(self, num_envs: int, lower, upper, initial_pos, action_index, abdomen_local, leg_tip_local,
                 num_bodies: int = 93):
        self.num_envs = n = num_envs
        self.lower = torch.tensor(lower, dtype=torch.float32)
        self.upper = torch.tensor(upper, dtype=torch.float32)
        self.num_dof = d = len(lower)
        self.num_bodies = num_bodies
        self.action_idx_one = torch.tensor(sorted(action_index), dtype=torch.long).view(-1, 1)  # fly.py:288-290
        self.num_act = a = self.action_idx_one.shape[0]
        self.num_obs = 19 + 3 * a                                                               # fly.py:32
        env_off = torch.arange(n, dtype=torch.long)
        self.action_indexes = (env_off.view(n, 1) * d + self.action_idx_one.view(1, a)).reshape(n * a, 1)
        self.abdomen_ids = (env_off.view(n, 1) * num_bodies +
                            torch.tensor(sorted(abdomen_local)).view(1, -1)).reshape(-1)         # fly.py:303-314
        self.leg_tip_ids = (env_off.view(n, 1) * num_bodies +
                            torch.tensor(sorted(leg_tip_local)).view(1, -1)).reshape(-1)
        one = torch.zeros(d, 2)
        one[:, 0] = torch.tensor(initial_pos, dtype=torch.float32)
        self.initial_dofs_one = one                                                              # fly.py:318-336
        self.initial_dofs = one.repeat(n, 1)
        origin = torch.zeros(n, 13)
        origin[:, 2] = STARTING_HEIGHT
        origin[:, 6] = 1.0
        self.origin_root = origin                                                                # fly.py:338-348
        self.targets = torch.tensor(TARGET).repeat(n, 1)                                         # fly.py:134
        self.inv_start_rot = im.quat_conjugate(torch.tensor([0.0, 0.0, 0.0, 1.0])).repeat(n, 1)  # fly.py:129
        self.basis0 = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)                                 # fly.py:127,131
        self.basis1 = torch.tensor(im.get_axis_params(1.0, 2), dtype=torch.float32).repeat(n, 1)  # fly.py:125,132


class EnvState:
    """Mutable buffers of the reference env (fly.py:169-179, 121-135) + the three Isaac Gym state views."""

    def __init__(self, t: EnvTables):
        n = t.num_envs
        self.obs_buf = torch.zeros(n, t.num_obs)
        self.reward_buf = torch.zeros(n)
        self.reset_buf = torch.ones(n, dtype=torch.long)
        self.progress_buf = torch.zeros(n, dtype=torch.long)
        self.potentials = torch.tensor([-1000.0 / DT]).repeat(n)
        self.prev_potentials = self.potentials.clone()
        self.up_vec = t.basis1.clone()
        self.heading_vec = t.basis0.clone()
        self.root = torch.zeros(n, 13)
        self.dof_states = torch.zeros(n, 2 * t.num_dof)
        self.force = torch.zeros(n * t.num_bodies, 3)
        self.actions = torch.zeros(n * t.num_act, 1)


def scale_actions(t: EnvTables, actions: Tensor) -> Tuple[Tensor, Tensor]:
    """reference fly.py:626-651: policy output in [-1,1] -> joint range; scatter into the DOF target vector.

    Returns (actions_scaled (N*A,1), dof position targets (N*D,))."""
    n = t.num_envs
    targets = torch.clone(t.initial_dofs)
    lo = t.lower[t.action_idx_one].view(1, -1).repeat(n, 1)
    hi = t.upper[t.action_idx_one].view(1, -1).repeat(n, 1)
    scaled = im.scale(actions, lo, hi).view(n * t.num_act, 1)
    targets[..., 0][t.action_indexes] = scaled
    return scaled.clone(), targets[..., 0].contiguous()


def apply_reset(t: EnvTables, s: EnvState) -> Tensor:
    """reference fly.py:446-480.  Returns the int32 env-id list handed to the two gym ``_indexed`` setters."""
    env_ids = s.reset_buf.nonzero(as_tuple=False).squeeze(-1)
    if len(env_ids) == 0:
        return env_ids.to(torch.int32)
    d = t.num_dof
    dof = s.dof_states.view(t.num_envs, d, 2)
    dof[env_ids, :, 0] = t.initial_dofs_one[..., 0]
    dof[env_ids, :, 1] = t.initial_dofs_one[..., 1]
    s.root[env_ids, :] = t.origin_root[env_ids, :]
    to_target = t.targets[env_ids] - t.origin_root[env_ids, 0:3]
    to_target[:, 2] = 0.0
    s.prev_potentials[env_ids] = -torch.norm(to_target, p=2, dim=-1) / DT
    s.potentials[env_ids] = s.prev_potentials[env_ids].clone()
    s.reset_buf[env_ids] = 0
    s.progress_buf[env_ids] = 0
    return env_ids.to(torch.int32)


def post_physics(t: EnvTables, s: EnvState, walking: bool = False,
                 energy_cost_scale: float = ENERGY_COST_SCALE, walking_leg_term: bool = True) -> None:
    """reference fly.py:676-681 → get_obs (397-411), progress += 1 (678), get_reward (413-443).
    ``energy_cost_scale``: 0.005 (fly.py:47) or 1 for the low-gravity variant (flyLowGrav.py:44)."""
    n, d = t.num_envs, t.num_dof
    dof = s.dof_states.view(n, d, 2)
    obs, pot, prev, up, head = observations(
        s.root, t.targets, s.potentials, t.inv_start_rot, dof[..., 0], dof[..., 1], t.lower, t.upper,
        DOF_VEL_SCALE, s.actions.view(n, -1), DT, t.basis0, t.basis1, t.action_idx_one, s.force, t.leg_tip_ids)
    s.obs_buf[:], s.potentials[:], s.prev_potentials[:], s.up_vec[:], s.heading_vec[:] = obs, pot, prev, up, head
    s.progress_buf += 1
    s.progress_buf[s.progress_buf == 0] = 1                                                      # fly.py:415-416
    rew, reset = reward_reset(
        s.obs_buf, s.reset_buf, s.progress_buf, s.actions.view(n, -1), s.potentials, s.prev_potentials,
        s.root[:, 3:7], s.force, t.abdomen_ids, t.leg_tip_ids, t.action_idx_one, t.upper, t.lower, t.num_act,
        UP_WEIGHT, HEADING_WEIGHT, ACTIONS_COST_SCALE, energy_cost_scale, JOINTS_AT_LIMIT_COST_SCALE,
        TERMINATION_HEIGHT, TERMINATION_HEIGHT_UP, DEATH_COST, MAX_EPISODE_LENGTH, walking, walking_leg_term)
    s.reward_buf[:], s.reset_buf[:] = rew, reset


ENERGY_COST_SCALE_LOWGRAV = 1.0   # flyLowGrav.py:44


def env_step(t: EnvTables, s: EnvState, actions: Tensor, simulate, variant: str = "standard",
             walking: bool = False) -> Tuple[Tensor, Tensor]:
    """One ``Fly.step(actions)`` around a caller-supplied physics step ``simulate(s)`` (which refreshes ``s.root``,
    ``s.dof_states``, ``s.force`` the way gym.simulate + refresh_* do).

    ``variant="standard"``: reference fly.py:624-681 — scale, set targets, ``reset()``, simulate, obs / reward.
    ``variant="lowgrav"``: reference flyLowGrav.py:617-683 — scale, set targets, simulate, THEN ``reset()``
    (flyLowGrav.py:657-663), obs / reward with ``energy_cost_scale = 1`` (flyLowGrav.py:44): the reset rows written
    into the state tensors are what the observation of the same step reads.
    Returns (dof position targets, int32 reset ids)."""
    scaled, targets = scale_actions(t, actions)
    s.actions = scaled
    if variant == "standard":
        ids = apply_reset(t, s)
        simulate(s)
        post_physics(t, s, walking)
    elif variant == "lowgrav":
        simulate(s)
        ids = apply_reset(t, s)
        post_physics(t, s, walking, ENERGY_COST_SCALE_LOWGRAV)
    else:
        raise ValueError(f"variant {variant!r}")
    return targets, ids


# --------------------------------------------------------------------------- actor-critic net
PARAM_ORDER = [
    "shared_net.0.weight", "shared_net.0.bias", "shared_net.2.weight", "shared_net.2.bias",
    "to_mean.0.weight", "to_mean.0.bias", "to_mean.2.weight", "to_mean.2.bias",
    "to_value.0.weight", "to_value.0.bias", "to_value.2.weight", "to_value.2.bias",
]


def init_params(num_obs: int, num_act: int, seed: int = 0) -> Dict[str, Tensor]:
    """``torch.nn.Linear`` default init in the reference's module construction order (ppo.py:18-38)."""
    g = torch.Generator().manual_seed(seed)
    shapes = [(256, num_obs), (128, 256), (64, 128), (num_act, 64), (64, 128), (1, 64)]
    names = ["shared_net.0", "shared_net.2", "to_mean.0", "to_mean.2", "to_value.0", "to_value.2"]
    out: Dict[str, Tensor] = {}
    for name, (fo, fi) in zip(names, shapes):
        bound = 1.0 / math.sqrt(fi)
        out[name + ".weight"] = (torch.rand(fo, fi, generator=g) * 2 - 1) * bound
        out[name + ".bias"] = (torch.rand(fo, generator=g) * 2 - 1) * bound
    return out


def trunk(p: Dict[str, Tensor], x: Tensor) -> Tensor:
    h = F.elu(F.linear(x, p["shared_net.0.weight"], p["shared_net.0.bias"]))
    return F.elu(F.linear(h, p["shared_net.2.weight"], p["shared_net.2.bias"]))


def pi(p: Dict[str, Tensor], x: Tensor) -> Tensor:
    """reference ppo.py:88-94 + 26-31: ELU after every layer, including the last."""
    h = trunk(p, x)
    h = F.elu(F.linear(h, p["to_mean.0.weight"], p["to_mean.0.bias"]))
    return F.elu(F.linear(h, p["to_mean.2.weight"], p["to_mean.2.bias"]))


def v(p: Dict[str, Tensor], x: Tensor) -> Tensor:
    """reference ppo.py:96-102 + 34-38."""
    h = trunk(p, x)
    h = F.elu(F.linear(h, p["to_value.0.weight"], p["to_value.0.bias"]))
    return F.linear(h, p["to_value.2.weight"], p["to_value.2.bias"])


def gaussian_logp(mu: Tensor, x: Tensor, var: Tensor) -> Tensor:
    """Closed form of ``MultivariateNormal(mu, scale_tril=cholesky(diag(var))).log_prob(x)`` (ppo.py:186-189,
    215-219), written in torch.distributions' own order: -0.5*(k*log(2π) + M) - Σ log L_ii, M = Σ ((x-μ)/σ)²."""
    std = var.sqrt()
    diff = (x - mu) / std
    m = (diff * diff).sum(-1)
    half_log_det = std.log().sum(-1)
    return -0.5 * (mu.shape[-1] * math.log(2 * math.pi) + m) - half_log_det


def act(p: Dict[str, Tensor], obs: Tensor, var: Tensor, noise: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """reference ppo.py:213-220 with the Gaussian draw made explicit: a = μ + sqrt(var)·ε.

    Returns (clipped action, log-prob of the UNclipped sample, μ)."""
    with torch.no_grad():
        mu = pi(p, obs)
        sample = mu + var.sqrt() * noise
        logp = gaussian_logp(mu, sample, var)
        return sample.clip(-1, 1), logp, mu


def anneal_var(var: Tensor) -> Tensor:
    """reference ppo.py:236-237."""
    return torch.max(VAR_FLOOR * torch.ones_like(var), var - VAR_DECAY)


def rollout_shape(num_envs: int) -> Tuple[int, int]:
    """reference ppo.py:118-122: (mini_chunk_size, rollout_size)."""
    mc = MINI_BATCH_SIZE // num_envs
    return mc, mc * CHUNK_NUMBER


def make_data(p: Dict[str, Tensor], all_obs: Tensor, all_next_obs: Tensor, all_reward: Tensor,
              done_last: Tensor) -> Tuple[Tensor, Tensor]:
    """reference ppo.py:157-171.  ``done_last`` is the (N,1) mask of the LAST step (the ppo.py:230 quirk),
    broadcast over T; the recursion carries no done mask.  Returns (target (T,N,1), advantage (T,N,1))."""
    with torch.no_grad():
        target = all_reward + GAMMA * v(p, all_next_obs) * done_last
        delta = target - v(p, all_obs)
    adv = torch.zeros_like(delta)
    run = 0.0
    for t in range(delta.shape[0] - 1, -1, -1):
        run = GAMMA * LMBDA * run + delta[t]
        adv[t] = run
    return target, adv


def minibatch_loss(p: Dict[str, Tensor], obs: Tensor, action: Tensor, old_logp: Tensor, target: Tensor,
                   adv: Tensor, var: Tensor) -> Tensor:
    """reference ppo.py:184-194 then ``loss.mean()`` (197): scalar SmoothL1 mean added to every surrogate row."""
    logp = gaussian_logp(pi(p, obs), action, var)
    ratio = torch.exp(logp - old_logp).unsqueeze(-1)
    surr = torch.min(ratio * adv, torch.clamp(ratio, 1 - CLIP, 1 + CLIP) * adv)
    loss = -surr + F.smooth_l1_loss(v(p, obs), target)
    return loss.mean()


class AdamState:
    """torch.optim.Adam defaults (ppo.py:155): lr 1e-3, betas (0.9, 0.999), eps 1e-8, no weight decay."""

    def __init__(self, p: Dict[str, Tensor]):
        self.step = 0
        self.m = {k: torch.zeros_like(x) for k, x in p.items()}
        self.v = {k: torch.zeros_like(x) for k, x in p.items()}


def clip_and_adam(p: Dict[str, Tensor], grads: Dict[str, Tensor], st: AdamState, lr: float = LR) -> float:
    """``clip_grad_norm_(params, 1.0)`` (ppo.py:198) then ``Adam.step`` (199), in torch's operation order."""
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g, 2) for g in grads.values()]), 2)
    coef = torch.clamp(1.0 / (total + 1e-6), max=1.0)
    st.step += 1
    b1, b2, eps = 0.9, 0.999, 1e-8
    bc1 = 1 - b1 ** st.step
    bc2 = 1 - b2 ** st.step
    for k in p:
        g = grads[k] * coef
        st.m[k].lerp_(g, 1 - b1)
        st.v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (st.v[k].sqrt() / math.sqrt(bc2)).add_(eps)
        p[k].addcdiv_(st.m[k], denom, value=-(lr / bc1))
    return float(total)


def minibatch_step(p: Dict[str, Tensor], st: AdamState, obs: Tensor, action: Tensor, old_logp: Tensor,
                   target: Tensor, adv: Tensor, var: Tensor,
                   grad_hook=None) -> Tuple[float, float, Dict[str, Tensor]]:
    """One optimiser step of reference ppo.py:183-199.  ``grad_hook(grads)`` is where a data-parallel run
    averages gradients (before the clip).  Returns (loss, grad-norm before clip, grads)."""
    leaf = {k: x.detach().clone().requires_grad_(True) for k, x in p.items()}
    loss = minibatch_loss(leaf, obs, action, old_logp, target, adv, var)
    loss.backward()
    grads = {k: leaf[k].grad for k in p}
    if grad_hook is not None:
        grad_hook(grads)
    norm = clip_and_adam(p, grads, st)
    return float(loss), norm, grads


def update(p: Dict[str, Tensor], st: AdamState, all_obs: Tensor, all_acts: Tensor, all_logp: Tensor,
           all_next_obs: Tensor, all_reward: Tensor, done_last: Tensor, var: Tensor, mini_chunk: int,
           epochs: int = EPOCHS, grad_hook=None) -> List[float]:
    """reference ppo.py:173-202: 5 epochs × (T/mc − 1) unshuffled time-contiguous minibatches — the loop
    ``range(mc, T, mc)`` never visits the last chunk."""
    target, adv = make_data(p, all_obs, all_next_obs, all_reward, done_last)
    losses = []
    t_total = all_obs.shape[0]
    for _ in range(epochs):
        k = 0
        for j in range(mini_chunk, t_total, mini_chunk):
            loss, _, _ = minibatch_step(p, st, all_obs[k:j], all_acts[k:j], all_logp[k:j], target[k:j], adv[k:j],
                                        var, grad_hook)
            losses.append(loss)
            k = j
    return losses


# --------------------------------------------------------------------------- synthetic state source
def synth_state(t: EnvTables, seed: int, device: str = "cpu") -> Dict[str, Tensor]:
    """Seeded synthetic Isaac Gym state at the named shapes (SURVEY.md §8d value distributions): exercises every
    termination branch (low / high torso, flipped orientation, abdomen contact, episode timeout)."""
    g = torch.Generator().manual_seed(seed)
    n, d, b = t.num_envs, t.num_dof, t.num_bodies

    def U(*shape):
        return torch.rand(*shape, generator=g)

    def Nrm(*shape):
        return torch.randn(*shape, generator=g)

    root = torch.zeros(n, 13)
    root[:, 0:2] = U(n, 2) * 10 - 5
    sel = U(n)
    z = 2.0 + 0.3 * Nrm(n)
    z = torch.where(sel < 0.10, 0.5 + 0.6 * U(n), z)
    z = torch.where((sel >= 0.10) & (sel < 0.15), 1.1 + 0.3 * U(n), z)
    z = torch.where((sel >= 0.15) & (sel < 0.20), 6.0 + U(n), z)
    root[:, 2] = z
    kind = U(n)
    yaw = U(n) * 2 * math.pi - math.pi
    tilt = 0.05 * Nrm(n, 2)
    q_yaw = torch.stack([tilt[:, 0], tilt[:, 1], torch.sin(yaw / 2), torch.cos(yaw / 2)], dim=-1)
    q_rand = Nrm(n, 4)
    q_flip = torch.stack([torch.cos(yaw / 2), torch.sin(yaw / 2), 0.3 * Nrm(n), 0.3 * Nrm(n)], dim=-1)
    q = torch.where((kind < 0.7).unsqueeze(-1), q_yaw, torch.where((kind < 0.9).unsqueeze(-1), q_rand, q_flip))
    root[:, 3:7] = q / q.norm(dim=-1, keepdim=True)
    root[:, 7:13] = Nrm(n, 6)

    dof = torch.zeros(n, d, 2)
    dof[..., 0] = t.lower + (t.upper - t.lower) * U(n, d)
    dof[..., 1] = 2.0 * Nrm(n, d)

    force = Nrm(n * b, 3)
    force = torch.where((U(n * b) < 0.7).unsqueeze(-1), torch.zeros_like(force), force)
    # keep most abdomens off the ground so that not every env dies
    ab = force[t.abdomen_ids].view(n, -1, 3)
    ab = torch.where((U(n) < 0.85).view(n, 1, 1), torch.zeros_like(ab), ab)
    force[t.abdomen_ids] = ab.view(-1, 3)

    progress = torch.randint(1, 1601, (n,), generator=g)
    actions = (0.6 * Nrm(n, t.num_act)).clip(-1, 1)
    out = {"root": root, "dof_states": dof.view(n, 2 * d).contiguous(), "force": force,
           "progress": progress, "actions": actions}
    return {k: x.to(device) for k, x in out.items()}
