"""CPU arm of bench.py: the reference's OWN code for the hot path, timed on the host cores.

TEST / MEASUREMENT INFRASTRUCTURE (see oracle/__init__.py) — never on the product path.

What runs is the reference's unmodified `ppo.PPO` (`run`, `make_data`, `update`: reference ppo.py:105-264) driving an
env object behind the same duck-typed surface `Fly` offers (ppo.py:110-112, 210-211, 223, 228-230), whose `step` calls
the reference's unmodified `fly.compute_fly_observations` / `fly.compute_fly_reward2` (fly.py:685-805; for the walking
configuration the same function with its commented-out total switched back on, ref_harness.load_walking_reward) — imported from
`/root/reference`, or from the byte-for-byte staged copy `oracle/_ref` on the GPU box (oracle/make_ref.py,
oracle/ref_harness.py).  The ~40 lines of `Fly.step` / `Fly.reset` tensor math that cannot be imported (they sit between
gymapi calls, fly.py:626-657, 446-480) are the oracle's line-by-line restatement (fly_oracle.scale_actions /
apply_reset); the synthetic Isaac Gym state comes from the same seeded generator as the GPU arm.  `kind = "reference"`.

When the reference's files are not present the same measurements run on the oracle port alone (`kind = "port"`).

One training iteration of the reference is far too long to repeat K times on host cores (75 optimiser steps of
32 768 samples), so each piece is timed on a bounded sample and the iteration is composed from them:
    t_iter = T * t_env_step + t_make_data(T) + n_minibatches * t_minibatch
`t_env_step`: median `PPO.run()` call (act + env.step + buffer writes, ppo.py:204-237); `t_make_data`: one
`PPO.make_data()` over the (T, N) buffers (ppo.py:157-171); `t_minibatch`: the reference's `PPO.update()` run on a
rollout truncated to 2 mini-chunks (5 epochs x 1 minibatch of the TRUE size), make_data subtracted, divided by 5.
"""
from __future__ import annotations

import contextlib
import io
import os
import statistics
import time
import types
from typing import Dict

import torch

from . import fly_oracle as fo
from . import ref_harness


def _make_env_class(fly_mod, t: fo.EnvTables, states, walking: bool):
    """Env behind the surface ppo.py uses.  `fly_mod`: the reference's fly module (its jit functions) or None (port)."""
    n = t.num_envs
    reward_fn = None
    if fly_mod is not None:     # the live total is the standing one (fly.py:750); walking = its commented-out lines, run
        reward_fn = ref_harness.load_walking_reward() if walking else fly_mod.compute_fly_reward2

    class SynthFly:
        def __init__(self, args):
            self.num_act, self.num_obs, self.end = t.num_act, t.num_obs, False
            self.s = fo.EnvState(t)
            self.obs_buf, self.reward_buf, self.reset_buf = self.s.obs_buf, self.s.reward_buf, self.s.reset_buf
            self.k = 0

        def step(self, actions):
            s = self.s
            s.actions, _ = fo.scale_actions(t, actions)                         # fly.py:626-657
            fo.apply_reset(t, s)                                                # fly.py:446-480
            st = states[self.k % len(states)]
            self.k += 1
            s.root, s.dof_states, s.force = st["root"], st["dof_states"], st["force"]   # gym.simulate + refresh_*
            if fly_mod is None:
                fo.post_physics(t, s, walking=walking)
            else:
                dof = s.dof_states.view(n, t.num_dof, 2)
                s.obs_buf[:], s.potentials[:], s.prev_potentials[:], s.up_vec[:], s.heading_vec[:] = \
                    fly_mod.compute_fly_observations(
                        s.obs_buf, s.root, t.targets, s.potentials, t.inv_start_rot, dof[..., 0], dof[..., 1], t.lower,
                        t.upper, fo.DOF_VEL_SCALE, s.actions.view(n, -1), fo.DT, t.basis0, t.basis1, 2,
                        t.action_idx_one, s.force, t.leg_tip_ids, n)            # fly.py:405-411
                s.progress_buf += 1                                             # fly.py:678
                s.progress_buf[s.progress_buf == 0] = 1                         # fly.py:415-416
                s.reward_buf[:], s.reset_buf[:] = reward_fn(
                    s.obs_buf, s.reset_buf, s.progress_buf, s.actions.view(n, -1), fo.UP_WEIGHT, fo.HEADING_WEIGHT,
                    s.potentials, s.prev_potentials, fo.ACTIONS_COST_SCALE, fo.ENERGY_COST_SCALE,
                    fo.JOINTS_AT_LIMIT_COST_SCALE, fo.TERMINATION_HEIGHT, fo.TERMINATION_HEIGHT_UP, fo.DEATH_COST,
                    fo.MAX_EPISODE_LENGTH, t.action_idx_one, t.upper, t.lower, n, s.root[:, 3:7], t.num_act, s.force,
                    t.abdomen_ids, t.leg_tip_ids)                               # fly.py:418-443
            self.obs_buf, self.reward_buf, self.reset_buf = s.obs_buf, s.reward_buf, s.reset_buf

        def generate_video(self):
            pass

        def exit(self):
            pass

    return SynthFly


class _PortAgent:
    """The oracle port behind the same three calls (used only when the reference's files are absent)."""

    def __init__(self, env, n, horizon, mc):
        self.env, self.n, self.rollout_size, self.mini_chunk_size = env, n, horizon, mc
        self.p = fo.init_params(env.num_obs, env.num_act, seed=0)
        self.st = fo.AdamState(self.p)
        self.var = torch.full((env.num_act,), 0.2)
        self.gen = torch.Generator().manual_seed(0)
        o, a = env.num_obs, env.num_act
        self.all_obs = torch.zeros(horizon, n, o)
        self.all_next_obs = torch.zeros(horizon, n, o)
        self.all_acts = torch.zeros(horizon, n, a)
        self.all_reward = torch.zeros(horizon, n, 1)
        self.all_log_prob = torch.zeros(horizon, n)
        self.mini_batch_number = 0

    def run(self):
        i = self.mini_batch_number
        obs = self.env.obs_buf
        act, logp, _ = fo.act(self.p, obs, self.var, torch.randn(self.n, self.env.num_act, generator=self.gen))
        self.env.step(act)
        self.all_obs[i], self.all_acts[i], self.all_log_prob[i] = obs, act, logp
        self.all_next_obs[i] = self.env.obs_buf
        self.all_reward[i] = self.env.reward_buf.unsqueeze(-1)
        self.var = fo.anneal_var(self.var)
        self.mini_batch_number = (i + 1) % self.rollout_size

    def make_data(self):
        done = (1 - self.env.reset_buf).unsqueeze(-1)
        return fo.make_data(self.p, self.all_obs, self.all_next_obs, self.all_reward, done)

    def update(self):
        done = (1 - self.env.reset_buf).unsqueeze(-1)
        fo.update(self.p, self.st, self.all_obs, self.all_acts, self.all_log_prob, self.all_next_obs, self.all_reward,
                  done, self.var, self.mini_chunk_size)


def _truncate(agent, rows: int) -> Dict[str, torch.Tensor]:
    names = ["all_obs", "all_acts", "all_next_obs", "all_reward", "all_log_prob", "all_advantage"]
    saved = {k: getattr(agent, k) for k in names if hasattr(agent, k)}
    for k, v in saved.items():
        setattr(agent, k, v[:rows].clone())
    saved["rollout_size"] = agent.rollout_size
    agent.rollout_size = rows
    return saved


def _restore(agent, saved) -> None:
    for k, v in saved.items():
        setattr(agent, k, v)


class ReferenceRunner:
    """Builds the CPU arm for one workload and times its pieces.  `threads`: torch intra-op threads."""

    def __init__(self, tables: fo.EnvTables, horizon: int, mini_chunk: int, walking: bool, threads: int, seed: int = 1234):
        torch.set_num_threads(max(1, threads))
        self.t, self.horizon, self.mc, self.threads = tables, horizon, mini_chunk, threads
        n = tables.num_envs
        states = [fo.synth_state(tables, seed + 7919 * i) for i in range(2)]
        self.kind = "reference" if ref_harness.available() else "port"
        sink = io.StringIO()
        if self.kind == "reference":
            fly_mod = ref_harness.load_fly()
            env_cls = _make_env_class(fly_mod, tables, states, walking)
            ppo = ref_harness.load_ppo(env_cls)
            args = types.SimpleNamespace(num_envs=n, sim_device="cpu", load=False, load_path=None, save=False,
                                         save_path=None, save_freq=10 ** 9, testing=False)
            torch.manual_seed(0)
            with contextlib.redirect_stdout(sink):
                agent = ppo.PPO(args)
            # the reference derives T from N (ppo.py:118-122); the bench config names T explicitly
            agent.rollout_size, agent.mini_chunk_size = horizon, mini_chunk
            o, a = tables.num_obs, tables.num_act
            agent.all_obs = torch.zeros(horizon, n, o)
            agent.all_next_obs = torch.zeros(horizon, n, o)
            agent.all_acts = torch.zeros(horizon, n, a)
            agent.all_reward = torch.zeros(horizon, n, 1)
            agent.all_done = torch.zeros(horizon, n, 1)
            agent.all_log_prob = torch.zeros(horizon, n)
            agent.all_advantage = torch.zeros(horizon, n, 1)
            agent.num_eval_freq = 10 ** 9       # no periodic print inside the timed calls
            self.agent = agent
        else:
            env = _make_env_class(None, tables, states, walking)(None)
            self.agent = _PortAgent(env, n, horizon, mini_chunk)
        self._sink = sink

    def _run_once(self):
        a = self.agent
        if a.mini_batch_number + 1 >= a.rollout_size:      # never let run() fall into update() here
            a.mini_batch_number = 0
        with contextlib.redirect_stdout(self._sink):
            a.run()

    def time_env_steps(self, steps: int, warmup: int):
        for _ in range(warmup):
            self._run_once()
        ts = []
        for _ in range(steps):
            t0 = time.perf_counter()
            self._run_once()
            ts.append(time.perf_counter() - t0)
        return statistics.median(ts), ts

    def time_make_data(self) -> float:
        best = float("inf")
        for _ in range(2):                       # the first call pays allocator / thread-pool warm-up
            t0 = time.perf_counter()
            self.agent.make_data()
            best = min(best, time.perf_counter() - t0)
        return best

    def time_minibatch(self) -> float:
        """Seconds per optimiser step at the true minibatch size, through the reference's own update()."""
        a = self.agent
        saved = _truncate(a, 2 * self.mc)
        best = float("inf")
        try:
            for _ in range(2):                               # the first update() pays autograd / allocator warm-up
                t0 = time.perf_counter()
                a.make_data()
                t_md = time.perf_counter() - t0
                t0 = time.perf_counter()
                with contextlib.redirect_stdout(self._sink):
                    a.update()                               # 5 epochs x 1 minibatch (+ its own make_data)
                t_up = time.perf_counter() - t0
                best = min(best, max(t_up - t_md, 1e-9) / 5.0)
        finally:
            _restore(a, saved)
        return best

    def iteration(self, env_steps: int, warmup: int, with_update: bool = True) -> dict:
        n, t, mc = self.t.num_envs, self.horizon, self.mc
        t_env, samples = self.time_env_steps(env_steps, warmup)
        n_mb = 5 * (t // mc - 1)
        out = {"kind": self.kind, "cores": self.threads, "t_env_step_s": t_env, "env_steps_per_s_rollout": n / t_env,
               "n_minibatches": n_mb}
        if with_update:
            t_md = self.time_make_data()
            t_mb = self.time_minibatch()
            t_iter = t * t_env + t_md + n_mb * t_mb
            out.update({"t_make_data_s": t_md, "t_minibatch_s": t_mb, "t_iteration_s": t_iter,
                        "env_steps_per_s": n * t / t_iter, "update_samples_per_s": n_mb * mc * n / (t_md + n_mb * t_mb)})
        return out
