"""`Fly` — drop-in for the reference env class behind the same duck-typed surface (reference fly.py:10-808).

What `ppo.py` / `trainer.py` use (reference ppo.py:110-112, 210-211, 223, 228-230, 276, 279): ``Fly(args)``,
``num_act``, ``num_obs``, ``obs_buf``, ``reward_buf``, ``reset_buf``, ``end``, ``step(actions)``,
``generate_video()``, ``exit()``; also kept: ``progress_buf, potentials, prev_potentials, up_vec, heading_vec,
actions (N*A,1), dof_pos, dof_vel, root_tensor, force_tensor, targets`` (fly.py:54, 88-100, 121-135, 636), with the
reference's dtypes (``reset_buf`` / ``progress_buf`` int64, ``reset_buf`` initialised to ones, fly.py:169-179).

``step`` is two launches around the physics step: ``flyb_env_pre_physics`` (fly.py:626-660) and
``flyb_env_post_physics`` (fly.py:676-681).  No host synchronisation happens inside (the reference has three per step).

``args.variant = "lowgrav"`` selects the reference's low-gravity twin (flyLowGrav.py): ``energy_cost_scale = 1``
(flyLowGrav.py:44) and ``reset()`` moved AFTER ``simulate()`` (flyLowGrav.py:657-663), i.e. scale + targets ->
simulate -> reset -> obs / reward, so the observation of a step already sees the reset rows.
"""
from __future__ import annotations

import ctypes as C
import types

import torch

from . import _lib, dp
from .state_source import SyntheticStateSource
from .tables import make_env_cfg


def _get(args, name, default):
    return getattr(args, name, default) if args is not None else default


class Fly:
    def __init__(self, args, state_source=None):
        self.args = args
        self.end = False
        self.lib = _lib.lib()                                   # raises if libflyb200.so is missing
        n = int(_get(args, "num_envs", 1000))
        self.device = torch.device(_get(args, "sim_device", "cuda:0"))
        if self.device.type != "cuda":
            raise _lib.FlybError(f"flyb200.Fly runs on CUDA devices only (sim_device={self.device}); no CPU fallback")
        torch.cuda.set_device(self.device)
        self.variant = str(_get(args, "variant", "standard"))
        self.reset_after_simulate = self.variant == "lowgrav"      # flyLowGrav.py:657-663
        self.cfg = make_env_cfg(n, asset=_get(args, "asset", "limited_18"),
                                full_actions=_get(args, "full_actions", None), walking=bool(_get(args, "walking", False)),
                                variant=self.variant)
        cfg = self.cfg
        self.num_envs, self.num_dof, self.num_act, self.num_bodies = n, cfg.num_dof, cfg.num_act, cfg.num_bodies
        self.num_obs = 19 + 3 * self.num_act                    # fly.py:32
        self.dt, self.up_axis_idx = cfg.dt, 2
        self.max_episode_length = int(cfg.max_episode_length)
        dev = self.device
        f32 = dict(dtype=torch.float32, device=dev)

        # fly.py:169-179
        self.obs_buf = torch.zeros((n, self.num_obs), **f32)
        self.reward_buf = torch.zeros(n, **f32)
        self.reset_buf = torch.ones(n, dtype=torch.long, device=dev)
        self.progress_buf = torch.zeros(n, dtype=torch.long, device=dev)
        # fly.py:121-135
        self.potentials = torch.full((n,), -1000.0 / cfg.dt, **f32)
        self.prev_potentials = self.potentials.clone()
        self.up_vec = torch.tensor(list(cfg.basis1), **f32).repeat(n, 1)
        self.heading_vec = torch.tensor(list(cfg.basis0), **f32).repeat(n, 1)
        self.basis_vec0, self.basis_vec1 = self.heading_vec.clone(), self.up_vec.clone()
        self.targets = torch.tensor(list(cfg.target), **f32).repeat(n, 1)
        self.inv_start_rot = torch.tensor(list(cfg.inv_start_rot), **f32).repeat(n, 1)
        self.dof_limits_lower = torch.tensor(list(cfg.lower)[:cfg.num_dof], **f32)
        self.dof_limits_upper = torch.tensor(list(cfg.upper)[:cfg.num_dof], **f32)
        self.action_indexes_one = torch.tensor(list(cfg.action_idx)[:cfg.num_act], dtype=torch.long,
                                               device=dev).view(-1, 1)
        self._actions_scaled = torch.zeros((n, self.num_act), **f32)   # `self.actions` (fly.py:636)
        self.dof_targets = torch.zeros(n * cfg.num_dof, **f32)         # what set_dof_position_target_tensor gets
        self.reset_ids = torch.zeros(n, dtype=torch.int32, device=dev)
        self.reset_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self.reward_sum = torch.zeros(1, **f32)
        self._scratch = torch.zeros(self.lib.flyb_env_scratch_bytes(n), dtype=torch.uint8, device=dev)

        self.state = state_source if state_source is not None else SyntheticStateSource(
            cfg, dev, pool=int(_get(args, "state_pool", 4)), seed=dp.rank_seed(int(_get(args, "seed", 1234))))
        self.render_count = 0

    # ---- reference attribute names for the Isaac Gym views (fly.py:88-100, 395)
    @property
    def root_tensor(self):
        return self.state.root

    @property
    def dof_states(self):
        return self.state.dof_state

    @property
    def force_tensor(self):
        return self.state.force

    @property
    def dof_pos(self):
        return self.state.dof_state.view(self.num_envs, self.num_dof, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.state.dof_state.view(self.num_envs, self.num_dof, 2)[..., 1]

    @property
    def actions(self):
        return self._actions_scaled.view(self.num_envs * self.num_act, 1)

    # ---- the two halves of step()
    def pre_physics(self, actions: torch.Tensor) -> None:
        """fly.py:626-660: scale + scatter to DOF targets, then reset() of the envs whose mask is set (the low-gravity
        variant leaves the reset to `reset()` after `simulate()`)."""
        if actions.shape != (self.num_envs, self.num_act) or actions.dtype != torch.float32:
            raise _lib.FlybError(f"actions must be float32 ({self.num_envs}, {self.num_act}); got "
                                 f"{actions.dtype} {tuple(actions.shape)}")
        p = _lib.ptr
        st = self.state
        _lib.check(self.lib.flyb_env_pre_physics(
            C.byref(self.cfg), p(actions), p(self._actions_scaled), p(self.dof_targets),
            None if self.reset_after_simulate else p(self.reset_buf),
            p(self.progress_buf), p(self.potentials), p(self.prev_potentials), p(st.root), p(st.dof_state),
            _lib.current_stream()), "flyb_env_pre_physics")
        st.set_dof_position_targets(self.dof_targets)
        if not self.reset_after_simulate:
            st.set_reset_ids(self.reset_ids, self.reset_count)

    def simulate(self) -> None:
        self.state.simulate()

    def post_physics(self, obs_copy=None, reward_copy=None, tail=None) -> None:
        """fly.py:676-681: get_obs, progress_buf += 1, get_reward — one launch; also compacts the reset ids.
        `tail` (a `_lib.StepTail`): end-of-step housekeeping of PPO.run that rides on this launch (ppo.py:236-237)."""
        p = _lib.ptr
        st = self.state
        _lib.check(self.lib.flyb_env_post_physics(
            C.byref(self.cfg), p(st.root), p(st.dof_state), p(st.force), p(self._actions_scaled),
            p(self.potentials), p(self.prev_potentials), p(self.progress_buf), p(self.reset_buf), p(self.obs_buf),
            p(self.reward_buf), p(self.up_vec), p(self.heading_vec), p(self.reset_ids), p(self.reset_count),
            p(self._scratch), p(obs_copy), p(reward_copy), p(self.reward_sum),
            C.byref(tail) if tail is not None else None, _lib.current_stream()),
            "flyb_env_post_physics")

    def step(self, actions: torch.Tensor) -> None:
        self.pre_physics(actions)
        self.simulate()
        if self.reset_after_simulate:       # flyLowGrav.py:657-663
            self.reset_async()
        self.render_count += 1
        self.post_physics()

    def reset_async(self) -> None:
        """Fly.reset() (fly.py:446-480) without the host round trip: flagged envs get the default DOF state and the
        origin root row, potentials re-initialised, mask / progress cleared — one stream-ordered call."""
        p = _lib.ptr
        st = self.state
        _lib.check(self.lib.flyb_env_reset(
            C.byref(self.cfg), p(self.reset_buf), p(self.progress_buf), p(self.potentials), p(self.prev_potentials),
            p(st.root), p(st.dof_state), _lib.current_stream()), "flyb_env_reset")
        st.set_reset_ids(self.reset_ids, self.reset_count)

    def reset(self) -> bool:
        """fly.py:446-480: resets the flagged envs and returns whether there were any (this call synchronises, like the
        reference's `len(env_ids)`; `step` never calls it — it uses the fused / asynchronous forms)."""
        any_flagged = bool(int(self.reset_buf.sum().item()) > 0)
        if any_flagged:
            self.reset_async()
        return any_flagged

    def generate_video(self) -> None:
        """Viewer / camera recording (fly.py:350-376, 487-615) needs a graphics stack: unavailable on this box."""

    def exit(self) -> None:
        pass


def make_args(**kw):
    """SimpleNamespace with the `args` fields trainer.py defines (reference trainer.py:6-20)."""
    base = dict(sim_device="cuda:0", compute_device_id=0, graphics_device_id=0, num_envs=1000, headless=True,
                save=False, save_path=None, save_freq=100, load=False, load_path=None, record=False,
                record_dir_name=None, time_steps_per_recorded_frame=2, testing=False)
    # beyond trainer.py: variant = "standard" | "lowgrav" (flyLowGrav.py), asset, full_actions, walking, horizon, seed,
    # state_pool, tensor_cores, verbose — all optional, read with defaults by Fly / PPO
    base.update(kw)
    return types.SimpleNamespace(**base)
