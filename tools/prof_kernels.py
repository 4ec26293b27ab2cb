#!/usr/bin/env python
"""Small driver for ncu captures: launches each hot-path kernel a few times at the bench workload (configs[1]).

    python tools/prof_kernels.py [act|post|step|vfwd|gae|grad|all] [--iters N] [--fp32]
"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import fly_bproject_b200 as fb  # noqa: E402
from fly_bproject_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("which", nargs="?", default="all")
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--envs", type=int, default=16384)
    ap.add_argument("--horizon", type=int, default=32)
    ap.add_argument("--fp32", action="store_true")
    a = ap.parse_args()
    dev = "cuda:0"
    torch.manual_seed(0)
    args = fb.make_args(num_envs=a.envs, sim_device=dev, horizon=a.horizon, verbose=False, state_pool=4, seed=1,
                        tensor_cores=not a.fp32)
    agent = fb.PPO(args)
    env = agent.env
    p = _lib.ptr
    for i in range(a.horizon):
        agent._step_kernels(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(name, fn):
        fn(0)
        torch.cuda.synchronize()
        e0.record()
        for i in range(a.iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        print(f"{name}: {e0.elapsed_time(e1) / a.iters * 1e3:.2f} us/launch", flush=True)

    def k_act(i):
        st = env.state
        _lib.check(agent.lib.flyb_policy_act(
            C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack), p(env.obs_buf), a.envs,
            p(agent.action_var), None, 1, i, None, p(agent.all_acts[i % a.horizon]),
            p(agent.all_log_prob[i % a.horizon]), None, None, C.byref(env.cfg), p(env._actions_scaled),
            p(env.dof_targets), p(env.reset_buf), p(env.progress_buf), p(env.potentials), p(env.prev_potentials),
            p(st.root), p(st.dof_state), _lib.current_stream()))

    def k_post(i):
        env.simulate()
        env.post_physics(obs_copy=agent._obs_ring[(i % a.horizon) + 1], reward_copy=agent.all_reward[i % a.horizon])

    def k_vfwd(i):
        _lib.check(agent.lib.flyb_net_forward(C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack),
                                              p(agent._obs_ring), (a.horizon + 1) * a.envs, None, p(agent._values),
                                              _lib.current_stream()))

    def k_gae(i):
        _lib.check(agent.lib.flyb_gae(p(agent.all_reward), p(agent._values), p(env.reset_buf), a.horizon, a.envs, 0.99,
                                      0.95, p(agent._target), p(agent.all_advantage), _lib.current_stream()))

    def k_grad(i):
        agent._minibatch(0, agent.mini_chunk_size)

    def k_step(i):
        agent._fused_step(1 + i % (a.horizon - 1))

    table = {"act": k_act, "post": k_post, "vfwd": k_vfwd, "gae": k_gae, "grad": k_grad}
    if agent._fused_steps:
        table["step"] = k_step
    for name, fn in table.items():
        if a.which in ("all", name):
            timed(name, fn)


if __name__ == "__main__":
    main()
