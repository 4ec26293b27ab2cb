"""`Net` and `PPO` — drop-ins for the reference agent behind the same surface (reference ppo.py:10-279).

`trainer.py` drives ``PPO(args)``, ``.run() -> bool``, ``.save(endofname="")``, ``.generate_video()``, ``.exit()``
(reference trainer.py:39-50); attribute names, buffer shapes / dtypes, hyper-parameters, the annealing schedule and
every quirk listed in SURVEY.md §7 are kept (trailing ELU on mu; log-prob of the UNclipped sample but the clipped
action stored; ``all_done`` = the last step's (N,1) mask; no done mask in the GAE recursion; the last mini-chunk never
trained on; the update uses the CURRENT annealed variance; scalar SmoothL1 mean added to every surrogate row).

All tensor math runs in libflyb200.so (csrc/policy_kernels.cu, gae_kernels.cu, ppo_kernels.cu).  PyTorch only owns
device memory, streams, CUDA graphs and the process group that carries the IPC handles / the parameter broadcast; the
data-parallel gradient sum itself is the library's own kernels over NVLink peer memory (dp.PeerExchange; NCCL only as
`FLYB_DP=nccl` or when peer mapping fails).
"""
from __future__ import annotations

import ctypes as C
import os
from collections import OrderedDict
from typing import Optional

import torch

from . import _lib, checkpoint, dp
from .env import Fly, _get

PARAM_NAMES = [
    "shared_net.0.weight", "shared_net.0.bias", "shared_net.2.weight", "shared_net.2.bias",
    "to_mean.0.weight", "to_mean.0.bias", "to_mean.2.weight", "to_mean.2.bias",
    "to_value.0.weight", "to_value.0.bias", "to_value.2.weight", "to_value.2.bias",
]


class Net:
    """Actor-critic parameters in ONE flat device buffer (FlybNetLayout) + the reference's `pi` / `v` (ppo.py:88-102).

    ``state_dict()`` / ``load_state_dict()`` speak the reference's checkpoint format (``torch.save(net.state_dict())``,
    ppo.py:266-273): the 12 tensors of ``nn.Sequential`` names above, so the 95 live-architecture ``.pth`` files load.
    """

    def __init__(self, num_obs: int, num_act: int, device="cuda:0", init: bool = True, tensor_cores: bool = True):
        """`tensor_cores`: run the GEMMs as 3xTF32 products on tcgen05 where a compiled shape covers the net
        (flyb_wpack_bytes > 0): num_obs in (64, 80] with num_act <= 32 — forward, act, fused env step and the PPO
        step — or anything else up to num_obs <= 160, num_act <= 48 — forward and act; the PPO step too when num_obs is
        in (144, 160], the 42-joint set.  False, or a shape outside both = the exact-fp32 CUDA-core kernels."""
        self.lib = _lib.lib()
        self.device = torch.device(device)
        self.layout = _lib.NetLayout()
        _lib.check(self.lib.flyb_net_layout(num_obs, num_act, C.byref(self.layout)), "flyb_net_layout")
        self.num_obs, self.num_act = num_obs, num_act
        self.flat = torch.zeros(self.layout.total, dtype=torch.float32, device=self.device)
        self._views = self.make_views(self.flat)
        wbytes = self.lib.flyb_wpack_bytes(C.byref(self.layout)) if tensor_cores else 0
        self.wpack = torch.zeros(wbytes, dtype=torch.uint8, device=self.device) if wbytes else None
        if init:
            self.reset_parameters()

    def shapes(self):
        o, a = self.num_obs, self.num_act
        return [(256, o), (256,), (128, 256), (128,), (64, 128), (64,), (a, 64), (a,), (64, 128), (64,), (1, 64), (1,)]

    def make_views(self, flat: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """Named views into a flat buffer of this layout (the first weight is a strided view over padded rows)."""
        lay, out = self.layout, OrderedDict()
        for i, (name, shp) in enumerate(zip(PARAM_NAMES, self.shapes())):
            off = lay.off[i]
            if i == 0:
                out[name] = flat[off:off + 256 * lay.k0_pad].view(256, lay.k0_pad)[:, :self.num_obs]
            else:
                n = 1
                for s in shp:
                    n *= s
                out[name] = flat[off:off + n].view(*shp)
        return out

    def reset_parameters(self) -> None:
        """nn.Linear default init drawn from torch's global RNG in the reference's construction order (ppo.py:18-38),
        so `torch.manual_seed(0)` (trainer.py:24) yields the reference's initial weights."""
        o, a = self.num_obs, self.num_act
        mods = [torch.nn.Linear(o, 256), torch.nn.Linear(256, 128), torch.nn.Linear(128, 64),
                torch.nn.Linear(64, a), torch.nn.Linear(128, 64), torch.nn.Linear(64, 1)]
        sd = OrderedDict()
        for prefix, m in zip(["shared_net.0", "shared_net.2", "to_mean.0", "to_mean.2", "to_value.0", "to_value.2"], mods):
            sd[prefix + ".weight"], sd[prefix + ".bias"] = m.weight.detach(), m.bias.detach()
        self.load_state_dict(sd)

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, v.detach().clone().contiguous()) for k, v in self._views.items())

    def load_state_dict(self, sd) -> str:
        missing = [k for k in PARAM_NAMES if k not in sd]
        unexpected = [k for k in sd if k not in PARAM_NAMES]
        if missing or unexpected:
            raise RuntimeError(f"Error(s) in loading state_dict for Net: missing {missing}, unexpected {unexpected}")
        with torch.no_grad():
            for k, v in self._views.items():
                src = sd[k]
                if tuple(src.shape) != tuple(v.shape):
                    raise RuntimeError(f"size mismatch for {k}: checkpoint {tuple(src.shape)} vs model {tuple(v.shape)}")
                v.copy_(src.to(device=self.device, dtype=torch.float32))
        self.repack()
        return "<All keys matched successfully>"

    def repack(self) -> None:
        """Refresh the tensor-core weight stream after `flat` changed (flyb_ppo_apply does it in-kernel order)."""
        if self.wpack is not None:
            _lib.check(self.lib.flyb_pack_weights(C.byref(self.layout), _lib.ptr(self.flat), _lib.ptr(self.wpack),
                                                  _lib.current_stream()), "flyb_pack_weights")

    def parameters(self):
        return list(self._views.values())

    def to(self, device):
        if torch.device(device) != self.device:
            raise _lib.FlybError("flyb200.Net lives on the CUDA device it was created on")
        return self

    def _forward(self, x: torch.Tensor, want_pi: bool, want_v: bool):
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self.num_obs)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        mu = torch.empty((rows, self.num_act), dtype=torch.float32, device=self.device) if want_pi else None
        val = torch.empty(rows, dtype=torch.float32, device=self.device) if want_v else None
        _lib.check(self.lib.flyb_net_forward(C.byref(self.layout), _lib.ptr(self.flat), _lib.ptr(self.wpack),
                                             _lib.ptr(x2), rows,
                                             _lib.ptr(mu), _lib.ptr(val), _lib.current_stream()), "flyb_net_forward")
        return (mu.view(*lead, self.num_act) if want_pi else None, val.view(*lead, 1) if want_v else None)

    def pi(self, x: torch.Tensor) -> torch.Tensor:
        return self._forward(x, True, False)[0]

    def v(self, x: torch.Tensor) -> torch.Tensor:
        return self._forward(x, False, True)[1]


class PPO:
    def __init__(self, args, env=None):
        self.args = args
        self.lib = _lib.lib()
        self.env = env if env is not None else Fly(args)
        self.device = torch.device(_get(args, "sim_device", "cuda:0"))
        n = int(args.num_envs)
        self.num_envs = n
        self.num_acts = self.env.num_act
        self.num_obs = self.env.num_obs
        # reference ppo.py:113-125
        self.epoch = 5
        self.lr = 0.001
        self.gamma = 0.99
        self.lmbda = 0.95
        self.clip = 0.2
        self.mini_batch_size = 40960
        self.chuck_number = 16
        horizon = _get(args, "horizon", None)
        if horizon is None:
            self.mini_chunk_size = self.mini_batch_size // n           # ppo.py:120 (0 for N > 40960: pass args.horizon)
            self.rollout_size = self.mini_chunk_size * self.chuck_number
        else:
            self.rollout_size = int(horizon)
            self.mini_chunk_size = max(1, self.rollout_size // self.chuck_number)
        if self.rollout_size <= 0:
            raise ValueError(f"rollout_size is 0 for num_envs={n} (reference ppo.py:120 breaks above 40960 envs); "
                             "set args.horizon")
        self.verbose = bool(_get(args, "verbose", True))
        if self.verbose:
            print("mini_chunk_size: ", self.mini_chunk_size)
            print("rollout_size: ", self.rollout_size)
        self.num_eval_freq = 100
        self.mini_batch_number = 0

        t, o, a, dev = self.rollout_size, self.num_obs, self.num_acts, self.device
        f32 = dict(dtype=torch.float32, device=dev)
        # ppo.py:132-138.  all_obs / all_next_obs are two views of one (T+1, N, O) ring: all_next_obs[t] IS
        # all_obs[t+1] in the reference too (ppo.py:210, 228), so the second buffer and value forward are not needed.
        self._obs_ring = torch.zeros((t + 1, n, o), **f32)
        self.all_obs = self._obs_ring[:t]
        self.all_next_obs = self._obs_ring[1:]
        self.all_acts = torch.zeros((t, n, a), **f32)
        self.all_reward = torch.zeros((t, n, 1), **f32)
        self.all_log_prob = torch.zeros((t, n), **f32)
        self.all_advantage = torch.zeros((t, n, 1), **f32)
        self._target = torch.zeros((t, n, 1), **f32)
        self._values = torch.zeros((t + 1, n), **f32)

        self.score = 0
        self.run_step = 0
        self.optim_step = 0

        self.net = Net(o, a, dev, tensor_cores=bool(_get(args, "tensor_cores", True)))
        resume_state = None
        if _get(args, "load", False):
            if self.verbose:
                print("loaded from: ", str(args.load_path))
            # ppo.py:147-149; a `.flyb.pt` sidecar next to the file (checkpoint.py) also brings back what the reference
            # loses on resume, unless args.resume is False.  Evaluation (args.testing: trainer.py --testing --load_path,
            # ppo.py:151) takes the weights only: it must act at action_var = 0.01 with fresh counters, not at the
            # saved training variance
            want_sidecar = bool(_get(args, "resume", True)) and not bool(_get(args, "testing", False))
            sd, resume_state = checkpoint.load_checkpoint(args.load_path, o, a, want_sidecar=want_sidecar)
            self.net.load_state_dict(sd)

        action_var = 0.01 if _get(args, "testing", False) else 0.2                    # ppo.py:151-152
        self.action_var = torch.full((a,), action_var, **f32)

        # torch.optim.Adam(net.parameters(), lr) state (ppo.py:155), flat like the parameters
        self.exp_avg = torch.zeros_like(self.net.flat)
        self.exp_avg_sq = torch.zeros_like(self.net.flat)
        self.adam_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.grad = torch.zeros_like(self.net.flat)
        self.loss = torch.zeros(1, **f32)
        self.grad_norm = torch.zeros(1, **f32)
        self.ppo_cfg = _lib.PpoCfg(self.clip, self.lr, 0.9, 0.999, 1e-8, 1.0)
        rows = self.mini_chunk_size * n
        ws_bytes = self.lib.flyb_ppo_workspace_bytes(C.byref(self.net.layout), rows)
        self._workspace = torch.zeros(ws_bytes + 256, dtype=torch.uint8, device=dev)
        self._ws_ptr = (self._workspace.data_ptr() + 255) // 256 * 256

        # data-parallel ranks draw different action noise from identical args (seed + rank)
        self.seed = dp.rank_seed(int(_get(args, "seed", 0)))
        self._rng_counter = torch.zeros(1, dtype=torch.int64, device=dev)   # Philox draw number, advanced on device
        self._fused_env = isinstance(self.env, Fly)
        self._foreign_reward_sum = torch.zeros((), **f32)
        testing = bool(_get(args, "testing", False))
        self._decay, self._floor = (0.0, 0.0) if testing else (0.00001, 0.01)             # ppo.py:236-237
        self._tail = _lib.StepTail(self.action_var.data_ptr(), a, self._decay, self._floor, self._rng_counter.data_ptr(), 1)
        # one launch per env step inside collect_rollout (flyb_env_step_fused): tensor-core path, standard step order
        # Default: where every 128-env tile has an SM of its own (N <= 128 x SMs).  Beyond that the fused CTAs run their
        # tiles' env and MLP phases back to back while the two-kernel schedule spreads the env phase over the whole
        # machine (measured, profiles/r02d_sweep_*: 34.9 vs 35.5 us per env step at 16384 envs, 82.9 vs 62.5 at 32768).
        fused_ok = bool(
            self._fused_env and self.net.wpack is not None and not self.env.reset_after_simulate
            and self.lib.flyb_env_step_fused_supported(C.byref(self.net.layout), C.byref(self.env.cfg)))
        want = _get(args, "fused_steps", None)
        if want is None:
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            want = (n + 127) // 128 <= sms
        self._fused_steps = bool(fused_ok and want)
        self._graphs = {}
        if resume_state is not None:
            self._restore(resume_state)

        # data-parallel: one process per GPU, the only collective is the gradient all-reduce
        self.rank, self.world_size = dp.world()
        self._peers = None                  # dp.PeerExchange: gradient sum by the library's own kernels over NVLink
        # how a data-parallel optimiser step exchanges gradients over peer memory (FLYB_DP_STEP):
        #   "fused"  flyb_ppo_step_dp: the reducing threads push self-validating words, await and sum the peers' ones; 4 launches
        #   "split"  flyb_ppo_grad (push + fence) + flyb_ppo_apply (norm kernel: flags, sum); 5 launches
        # FLYB_DP_TAIL=1 (with "fused"): the cooperative two-launch form of flyb_ppo_step_dp instead
        self._dp_step = os.environ.get("FLYB_DP_STEP", "fused").lower()
        if self._dp_step not in ("fused", "split"):
            raise ValueError(f"FLYB_DP_STEP={self._dp_step!r}: expected 'fused' or 'split'")
        if os.environ.get("FLYB_DP_TAIL", "0") == "1":
            self._dp_step = "fused"
        if self.world_size > 1:
            dp.broadcast_params_(self.net.flat, src=0)
            self.net.repack()
            self._peers = dp.make_exchange(self.net.layout.total, dev)

    # ------------------------------------------------------------------ reference-compatible views
    @property
    def all_done(self) -> torch.Tensor:
        """ppo.py:230 rebinds all_done to the LAST step's (N,1) int64 mask."""
        return (1 - self.env.reset_buf).unsqueeze(-1)

    @property
    def optim(self):
        return self

    def state(self):
        return {"exp_avg": self.net.make_views(self.exp_avg), "exp_avg_sq": self.net.make_views(self.exp_avg_sq),
                "step": self.adam_step}

    # ------------------------------------------------------------------ one env step (ppo.py:204-237)
    def _step_kernels(self, i: int, noise: Optional[torch.Tensor] = None) -> None:
        env, p = self.env, _lib.ptr
        lay = C.byref(self.net.layout)
        stream = _lib.current_stream()
        if self._fused_env:
            st = env.state
            _lib.check(self.lib.flyb_policy_act(
                lay, p(self.net.flat), p(self.net.wpack), p(env.obs_buf), self.num_envs, p(self.action_var), p(noise), self.seed, 0,
                p(self._rng_counter), p(self.all_acts[i]), p(self.all_log_prob[i]), None,
                p(self._obs_ring[0]) if i == 0 else None, C.byref(env.cfg), p(env._actions_scaled),
                p(env.dof_targets), None if env.reset_after_simulate else p(env.reset_buf), p(env.progress_buf),
                p(env.potentials), p(env.prev_potentials), p(st.root), p(st.dof_state), stream), "flyb_policy_act")
            st.set_dof_position_targets(env.dof_targets)
            if not env.reset_after_simulate:
                st.set_reset_ids(env.reset_ids, env.reset_count)
            env.simulate()
            if env.reset_after_simulate:        # low-gravity step order (flyLowGrav.py:657-663)
                env.reset_async()
            env.render_count += 1
            # ppo.py:236-237 (variance annealing, skipped in testing mode) and the Philox draw counter ride on the
            # step's last launch
            env.post_physics(obs_copy=self._obs_ring[i + 1], reward_copy=self.all_reward[i], tail=self._tail)
            return
        else:   # foreign env object behind the reference's duck-typed surface (ppo.py:210-229)
            obs = env.obs_buf.contiguous()
            _lib.check(self.lib.flyb_policy_act(
                lay, p(self.net.flat), p(self.net.wpack), p(obs), self.num_envs, p(self.action_var), p(noise), self.seed, 0,
                p(self._rng_counter), p(self.all_acts[i]), p(self.all_log_prob[i]), None, p(self._obs_ring[i]),
                None, None, None, None, None, None, None, None, None, stream), "flyb_policy_act")
            env.step(self.all_acts[i])
            self._obs_ring[i + 1].copy_(env.obs_buf)
            self.all_reward[i].copy_(env.reward_buf.unsqueeze(-1))
            self._foreign_reward_sum += env.reward_buf.sum()           # ppo.py:233, without its per-step .item()
        testing = bool(_get(self.args, "testing", False))
        # ppo.py:236-237 (skipped in testing mode); the same launch advances the Philox draw counter
        _lib.check(self.lib.flyb_anneal_var(p(self.action_var), self.num_acts, 0.0 if testing else 0.00001,
                                            0.0 if testing else 0.01, p(self._rng_counter), stream), "flyb_anneal_var")

    def run(self, noise: Optional[torch.Tensor] = None) -> bool:
        end = self.env.end
        self._step_kernels(self.mini_batch_number, noise)
        if self.mini_batch_number + 1 == self.rollout_size:
            if not _get(self.args, "testing", False):
                if self.verbose:
                    print("Training")
                self.update()
            self.mini_batch_number = 0
            if _get(self.args, "save", False) and self.optim_step % self.args.save_freq == 0 and self.optim_step != 0:
                self.save(str(self.optim_step))                                     # ppo.py:249-252
        else:
            self.mini_batch_number += 1
        if self.run_step % self.num_eval_freq == 0:                                 # ppo.py:257-260
            self._log()
        self.run_step += 1
        return end

    def _log(self) -> None:
        """The reference syncs every step for the score (ppo.py:233); here the sum lives on the device and is read
        once per `num_eval_freq` steps."""
        if self._fused_env:
            self.score = float(self.env.reward_sum.item()) / self.num_envs / self.num_eval_freq
            self.env.reward_sum.zero_()
        else:
            self.score = float(self._foreign_reward_sum.item()) / self.num_envs / self.num_eval_freq
            self._foreign_reward_sum.zero_()
        if self._peers is not None:
            self._peers.check()         # a timed-out exchange sets a sticky word instead of trapping (peer.cuh)
        if self.verbose:
            print('Steps: {:04d} | Opt Step: {:04d} | Reward {:.04f} | Action Var {:.04f}'.format(
                self.run_step, self.optim_step, self.score, self.action_var[0].item()))
        self.score = 0

    # ------------------------------------------------------------------ rollout with one launch per env step
    def _fused_step(self, i: int, noise: Optional[torch.Tensor] = None) -> None:
        """Launch i (1 <= i < T) of a fused rollout: obs + reward + reset of env step i-1, then act + pre-physics of env
        step i, for the same envs in the same CTA (flyb_env_step_fused)."""
        env, p, st = self.env, _lib.ptr, self.env.state
        io = _lib.StepIO(
            p(st.root), p(st.dof_state), p(st.force), p(env._actions_scaled), p(env.potentials), p(env.prev_potentials),
            p(env.progress_buf), p(env.reset_buf), None, p(self._obs_ring[i]), p(env.reward_buf),
            p(self.all_reward[i - 1]), p(env.reward_sum), p(env.up_vec), p(env.heading_vec), p(env.reset_ids),
            p(env.reset_count), p(env._scratch), p(self.action_var), p(noise), self.seed, i, p(self._rng_counter),
            p(self.all_acts[i]), p(self.all_log_prob[i]), p(env.dof_targets), i, self._decay, self._floor)
        _lib.check(self.lib.flyb_env_step_fused(C.byref(self.net.layout), p(self.net.flat), p(self.net.wpack),
                                                C.byref(env.cfg), C.byref(io), _lib.current_stream()),
                   "flyb_env_step_fused")
        st.set_dof_position_targets(env.dof_targets)
        st.set_reset_ids(env.reset_ids, env.reset_count)
        env.simulate()
        env.render_count += 1

    def _rollout_kernels(self, noise: Optional[torch.Tensor] = None) -> None:
        """The T env steps of one rollout.  Fused schedule (tensor-core path): act of step 0, T-1 fused launches, one
        closing obs + reward + reset launch whose tail applies the rollout's T annealing steps — T + 1 launches.  Same
        buffers, bit for bit, as T x `_step_kernels` (2T launches), which every other configuration uses.
        `noise` (T, N, A): explicit eps per step (parity tests)."""
        t = self.rollout_size
        if not self._fused_steps:
            for i in range(t):
                self._step_kernels(i, None if noise is None else noise[i])
            return
        env, p, st = self.env, _lib.ptr, self.env.state
        _lib.check(self.lib.flyb_policy_act(
            C.byref(self.net.layout), p(self.net.flat), p(self.net.wpack), p(env.obs_buf), self.num_envs,
            p(self.action_var), p(None if noise is None else noise[0]), self.seed, 0, p(self._rng_counter),
            p(self.all_acts[0]), p(self.all_log_prob[0]), None, p(self._obs_ring[0]), C.byref(env.cfg),
            p(env._actions_scaled), p(env.dof_targets), p(env.reset_buf), p(env.progress_buf), p(env.potentials),
            p(env.prev_potentials), p(st.root), p(st.dof_state), _lib.current_stream()), "flyb_policy_act")
        st.set_dof_position_targets(env.dof_targets)
        st.set_reset_ids(env.reset_ids, env.reset_count)
        env.simulate()
        env.render_count += 1
        for i in range(1, t):
            self._fused_step(i, None if noise is None else noise[i])
        tail = _lib.StepTail(self.action_var.data_ptr(), self.num_acts, self._decay, self._floor,
                             self._rng_counter.data_ptr(), t)
        env.post_physics(obs_copy=self._obs_ring[t], reward_copy=self.all_reward[t - 1], tail=tail)

    # ------------------------------------------------------------------ rollout as one CUDA graph
    def collect_rollout(self, use_graph: bool = True, noise: Optional[torch.Tensor] = None) -> None:
        """T env steps (act + pre-physics, state source, obs + reward + reset).  With `use_graph` the launches are
        captured once and replayed: the per-step pointers (rollout rows, state-pool entries) are baked in, the
        Philox draw number and action_var advance on the device."""
        t = self.rollout_size
        if not use_graph:
            self._rollout_kernels(noise)
            self.run_step += t
            return
        if noise is not None:
            raise _lib.FlybError("collect_rollout: explicit `noise` needs use_graph=False (the captured rollout draws "
                                 "from Philox; a replay could not pick up a new noise tensor)")
        g = self._graphs.get("rollout")
        if g is None:
            pool = len(getattr(self.env.state, "pool", [None]))
            if t % pool != 0:
                raise _lib.FlybError(f"graph capture needs rollout_size ({t}) to be a multiple of the state pool ({pool})")
            self._rollout_kernels()     # warm-up outside capture (also sets the kernels' smem attributes)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._rollout_kernels()
            self._graphs["rollout"] = g
            self.run_step += t
            return
        g.replay()
        self.run_step += t

    # ------------------------------------------------------------------ GAE (ppo.py:157-171)
    def make_data(self):
        t, n = self.rollout_size, self.num_envs
        p, stream = _lib.ptr, _lib.current_stream()
        _lib.check(self.lib.flyb_net_forward(C.byref(self.net.layout), p(self.net.flat), p(self.net.wpack),
                                             p(self._obs_ring),
                                             (t + 1) * n, None, p(self._values), stream), "flyb_net_forward")
        _lib.check(self.lib.flyb_gae(p(self.all_reward), p(self._values), p(self.env.reset_buf), t, n, self.gamma,
                                     self.lmbda, p(self._target), p(self.all_advantage), stream), "flyb_gae")
        return self.all_obs, self.all_acts, self.all_log_prob, self._target, self.all_advantage

    # ------------------------------------------------------------------ update (ppo.py:173-202)
    def _minibatch(self, k: int, j: int) -> None:
        p, stream = _lib.ptr, _lib.current_stream()
        rows = (j - k) * self.num_envs
        lay = C.byref(self.net.layout)
        if self.world_size == 1:     # nothing happens between the two halves: one call, two launches fewer
            _lib.check(self.lib.flyb_ppo_step(
                lay, C.byref(self.ppo_cfg), p(self.net.flat), p(self.net.wpack), p(self._obs_ring[k:j]),
                p(self.all_acts[k:j]), p(self.all_log_prob[k:j]), p(self._target[k:j]), p(self.all_advantage[k:j]),
                p(self.action_var), rows, p(self.grad), p(self.loss), p(self.exp_avg), p(self.exp_avg_sq),
                p(self.adam_step), p(self.grad_norm), self._ws_ptr, stream), "flyb_ppo_step")
            return
        peers = self._peers.byref() if self._peers is not None else None
        if peers is not None and self._dp_step == "fused":
            # chain, weight gradients, reduction WITH the exchange (push / await / rank-ordered sum over NVLink peer
            # memory), clip + Adam
            _lib.check(self.lib.flyb_ppo_step_dp(
                lay, C.byref(self.ppo_cfg), p(self.net.flat), p(self.net.wpack), p(self._obs_ring[k:j]),
                p(self.all_acts[k:j]), p(self.all_log_prob[k:j]), p(self._target[k:j]), p(self.all_advantage[k:j]),
                p(self.action_var), rows, p(self.grad), p(self.loss), p(self.exp_avg), p(self.exp_avg_sq),
                p(self.adam_step), 1.0 / self.world_size, p(self.grad_norm), self._ws_ptr, peers, stream),
                "flyb_ppo_step_dp")
            return
        _lib.check(self.lib.flyb_ppo_grad(
            lay, C.byref(self.ppo_cfg), p(self.net.flat), p(self.net.wpack), p(self._obs_ring[k:j]), p(self.all_acts[k:j]),
            p(self.all_log_prob[k:j]), p(self._target[k:j]), p(self.all_advantage[k:j]), p(self.action_var), rows,
            p(self.grad), p(self.loss), self._ws_ptr, peers, stream), "flyb_ppo_grad")
        if peers is None:
            grad_scale = dp.grad_all_reduce_(self.grad)      # SUM over ranks (NCCL); the mean is taken by grad_scale
        else:
            grad_scale = 1.0 / self.world_size               # the sum happens inside flyb_ppo_apply's first kernel
        _lib.check(self.lib.flyb_ppo_apply(
            lay, C.byref(self.ppo_cfg), p(self.net.flat), p(self.grad), p(self.exp_avg), p(self.exp_avg_sq),
            p(self.adam_step), grad_scale, p(self.grad_norm), self._ws_ptr, p(self.net.wpack), peers, stream),
            "flyb_ppo_apply")

    def _update_body(self) -> int:
        self.make_data()
        steps = 0
        for _ in range(self.epoch):
            k = 0
            for j in range(self.mini_chunk_size, self.rollout_size, self.mini_chunk_size):   # last chunk never visited
                self._minibatch(k, j)
                steps += 1
                k = j
        return steps

    def update(self, use_graph: bool = False) -> None:
        # a data-parallel run over NCCL keeps the all-reduce out of graph capture (kernels enqueued eagerly; the host
        # stays ahead of the device because one optimiser step is ~170 us of GPU work); with the peer exchange the
        # whole update is the library's own stream-ordered kernels and is captured like the single-GPU one
        if not use_graph or (self.world_size > 1 and self._peers is None):
            self.optim_step += self._update_body()
            return
        g = self._graphs.get("update")
        if g is None:
            self.optim_step += self._update_body()      # warm-up, eager
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                steps = self._update_body()
            self._graphs["update"] = (graph, steps)
            return
        graph, steps = g
        graph.replay()
        self.optim_step += steps

    def _restore(self, side: dict) -> None:
        """Adam moments, annealed variance, counters and the Philox draw number from a checkpoint sidecar."""
        if side["exp_avg"].numel() != self.exp_avg.numel():
            raise _lib.FlybError("checkpoint sidecar was written under a different parameter layout")
        self.exp_avg.copy_(side["exp_avg"])
        self.exp_avg_sq.copy_(side["exp_avg_sq"])
        self.adam_step.fill_(int(side["adam_step"]))
        self.action_var.copy_(side["action_var"])
        self._rng_counter.fill_(int(side["rng_counter"]))
        self.optim_step = int(side["optim_step"])
        self.run_step = int(side["run_step"])

    def save(self, endofname: str = "") -> None:
        """ppo.py:266-273: `save_path + endofname + ".pth"` in the reference's layout (loads in the reference), plus —
        unless args.save_sidecar is False — `<that>.flyb.pt` with Adam's moments and step, `action_var`, `optim_step`,
        `run_step` and the Philox draw counter, so that `args.load` resumes the run instead of restarting them."""
        if not _get(self.args, "save", False) or self.rank != 0:      # ranks hold identical parameters: rank 0 writes
            return
        path = self.args.save_path + endofname + ".pth"
        extras = None
        if _get(self.args, "save_sidecar", True):
            extras = {"exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "adam_step": int(self.adam_step.item()),
                      "action_var": self.action_var, "optim_step": self.optim_step, "run_step": self.run_step,
                      "rng_counter": int(self._rng_counter.item())}
        checkpoint.save_checkpoint(path, self.net.state_dict(), extras)

    def generate_video(self) -> None:
        self.env.generate_video()

    def exit(self) -> None:
        if self._peers is not None:
            peers, self._peers = self._peers, None
            try:
                peers.check()
            finally:
                peers.close()
        self.env.exit()
