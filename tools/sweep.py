#!/usr/bin/env python
"""Env-count sweep (BASELINE.json configs[4]): device time of the two per-step kernels at N = 1K .. 262144 envs, 18 DOF,
each as env-steps/s and as a fraction of its roofline (HBM for obs+reward+reset, 3xTF32 tensor pipe for act).

    python tools/sweep.py > profiles/<round>_sweep.json      (one JSON line per N; needs a GPU)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import fly_bproject_b200 as fb  # noqa: E402


def main():
    pk = bench.peaks()
    dev = "cuda:0"
    for n in (1024, 4096, 16384, 65536, 262144):
        pool = max(2, min(16, (400 << 20) // (n * 1320)))       # keep the pool above L2 where it can be
        args = fb.make_args(num_envs=n, sim_device=dev, horizon=16, verbose=False, state_pool=pool, seed=1)
        agent = fb.PPO(args)
        env = agent.env
        for i in range(16):
            agent._step_kernels(i)

        def k_env(i):
            env.simulate()
            env.post_physics(obs_copy=agent._obs_ring[(i % 16) + 1], reward_copy=agent.all_reward[i % 16])

        def k_step(i):
            agent._step_kernels(i % 16)

        t_env = bench.graph_time_ms(k_env, 4 * pool)
        t_step = bench.graph_time_ms(k_step, 16)
        t_act = t_step - t_env
        line = {
            "num_envs": n, "state_pool": pool, "pool_mb": env.state.pool_bytes / 1e6,
            "post_physics_us": t_env * 1e3, "act_us": t_act * 1e3, "env_step_us": t_step * 1e3,
            "env_steps_per_s": n / (t_step * 1e-3),
            "post_physics_gbs": bench.BYTES_ENV_KERNEL * n / t_env / 1e6,
            "post_physics_frac_hbm": bench.BYTES_ENV_KERNEL * n / t_env / 1e6 / pk["hbm"],
            "act_tflops": bench.FLOPS_ACT * n / t_act / 1e9,
            "act_frac_of_3xtf32_peak": bench.FLOPS_ACT * n / t_act / 1e9 / (pk["tf_sustained"] / 6.0),
        }
        print(json.dumps(line), flush=True)
        del agent, env
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
