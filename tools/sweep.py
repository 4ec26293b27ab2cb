#!/usr/bin/env python
"""BASELINE.json configs[4]: env-count x horizon sweep, 18 DOF, on 1 / 2 / 4 / 8 GPUs, with the host-core baseline.

    python tools/sweep.py [--envs 1024,4096,...] [--horizons 16,32,64,128] [--cpu] > profiles/<round>_sweep_n1.jsonl
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/sweep.py ...

One JSON line per (N per GPU, T) point: device time of the rollout (T fused env steps, one CUDA graph), of the PPO update
(value forward + GAE + 5 x 15 optimiser steps of N*T/16 samples, one graph, gradient exchange inside at G > 1) and of the
whole training iteration, each as a rate over all G GPUs (max over ranks), plus the rollout's tensor-roofline fraction.
`--cpu` adds the reference's own CPU code on the host cores (oracle/ref_bench.py) for the points with N <= 65536 and
T <= 32 (SURVEY.md §8d caps the CPU arm there; nothing is extrapolated beyond), rank 0 only.
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import fly_bproject_b200 as fb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", default="1024,4096,16384,65536,262144")
    ap.add_argument("--horizons", default="16,32,64,128")
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--fused", default="auto", choices=["auto", "on", "off"])
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":     # the banner goes to stdout at VERSION level
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    pk = bench.peaks()
    wm = bench.work_model(18, 18)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, k, reps=3):
        out = []
        for _ in range(reps):
            barrier()
            e0.record()
            for _ in range(k):
                fn()
            e1.record()
            barrier()
            out.append(rmax(e0.elapsed_time(e1)) / k)
        return statistics.median(out)

    for n in [int(x) for x in a.envs.split(",")]:
        for horizon in [int(x) for x in a.horizons.split(",")]:
            pool = max(2, min(16, (400 << 20) // (n * 1320)))       # keep the pool above L2 where it can be
            while horizon % pool:
                pool //= 2
            kw = {} if a.fused == "auto" else {"fused_steps": a.fused == "on"}
            args = fb.make_args(num_envs=n, sim_device=dev, horizon=horizon, verbose=False, state_pool=pool, seed=1, **kw)
            torch.manual_seed(0)
            agent = fb.PPO(args)
            mc = agent.mini_chunk_size
            n_mb = agent.epoch * (horizon // mc - 1)
            for _ in range(2):
                agent.collect_rollout(use_graph=True)
                agent.update(use_graph=True)
            k = max(2, min(20, int(2e8 // (n * horizon))))
            roll_ms = timed(lambda: agent.collect_rollout(use_graph=True), 2 * k)
            upd_ms = timed(lambda: agent.update(use_graph=True), k)
            it_ms = timed(lambda: (agent.collect_rollout(use_graph=True), agent.update(use_graph=True)), k)
            steps = n * horizon * world
            line = {
                "n_gpus": world, "num_envs_per_gpu": n, "horizon": horizon, "mini_chunk": mc, "state_pool": pool,
                "pool_mb": agent.env.state.pool_bytes / 1e6, "fused_steps": agent._fused_steps,
                "rollout_ms": roll_ms, "us_per_env_step": roll_ms * 1e3 / horizon, "env_steps_per_s_rollout": steps / roll_ms * 1e3,
                "rollout_frac_of_3xtf32_peak": wm["flops_act"] * n / (roll_ms / horizon) / 1e9 / (pk["tf_sustained"] / 6.0),
                "rollout_hbm_frac": wm["bytes_step"] * n / (roll_ms / horizon) / 1e6 / pk["hbm"],
                "update_ms": upd_ms, "update_samples_per_s": n_mb * mc * n * world / upd_ms * 1e3,
                "us_per_optimiser_step": upd_ms * 1e3 / n_mb,
                "update_frac_of_3xtf32_peak": wm["flops_update"] * mc * n * n_mb / upd_ms / 1e9 / (pk["tf_sustained"] / 6.0),
                "iteration_ms": it_ms, "env_steps_per_s": steps / it_ms * 1e3,
            }
            if a.cpu and rank == 0 and n <= 65536 and horizon <= 32:
                cfg = dict(asset="limited_18", full=True, walking=False, n=n, horizon=horizon)
                res = bench.cpu_runner(cfg, os.cpu_count() or 1).iteration(env_steps=8, warmup=4, with_update=True)
                line["cpu"] = {"kind": res["kind"], "cores": res["cores"], "env_steps_per_s": res["env_steps_per_s"],
                               "env_steps_per_s_rollout": res["env_steps_per_s_rollout"],
                               "update_samples_per_s": res["update_samples_per_s"]}
            if rank == 0:
                print(json.dumps(line), flush=True)
            if world > 1:
                agent.exit()
            del agent
            torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
