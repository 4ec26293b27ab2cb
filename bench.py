#!/usr/bin/env python
"""Benchmark of the fly_bProject hot path on B200 (contract: the task brief / DESIGN.md §7).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c0|c1|c2|c3|c3b] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One bench STEP = one pass of the whole hot path over the env batch = one TRAINING ITERATION of the reference's loop
(trainer.py:44-46 -> ppo.py:204-264): a T-step rollout — per env step [policy act + action scaling + reset bookkeeping]
-> [synthetic Isaac Gym state] -> [obs + reward + reset mask + id compaction], one fused launch per env step, replayed
as one CUDA graph — followed by the PPO update (value forward + GAE + 5 x (T/mc - 1) optimiser steps, one graph, with
the data-parallel gradient exchange inside it at N > 1).
`value` = env-steps/s through the whole iteration, whole job, inputs resident in HBM.  The two rates BASELINE.json's
metric names are reported beside it: `rollout` (fused env-steps/s: obs + reward + reset + act) and `update` (PPO update
samples/s).  The timed block of K iterations is repeated 5 times; the median block (max over ranks) is reported.
`e2e` = the same iteration through the public API (`PPO.run()` per env step, the T-th call runs the update) with HOST
state buffers: per env step an H2D copy of the three state tensors from pinned memory and a D2H read of what
`Fly.step` hands back (obs, reward, reset mask, compacted ids + count, DOF targets).

Default workload = BASELINE.json configs[1]: 18-DOF standing, 16384 envs per GPU, horizon 32.  `--config` selects the
other BASELINE configurations (c0: N=1000/T=640, c2: walking 8192 envs/GPU T=80, c3 / c3b: 42-DOF joint set N=32768
T=64 with 42 / 18 actuated joints).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "env_steps_per_sec"
UNIT = "env-steps/s"
MMA_KIND = "tcgen05 kind::tf32, 3xTF32 (hi*lo + lo*hi + hi*hi), fp32 accumulate in TMEM"

CONFIGS = {
    "c0": dict(label="configs[0]: NeuroMechFly 18-DOF standing, num_envs=1000, rollout 640 / mini-chunk 40 (trainer.py "
                     "defaults, ppo.py:118-122)", asset="limited_18", full=True, walking=False, n=1000, horizon=640, pool=16),
    "c1": dict(label="configs[1]: NeuroMechFly 18-DOF standing (A=D=18, O=73), num_envs=16384 per GPU, horizon 32",
               asset="limited_18", full=True, walking=False, n=16384, horizon=32, pool=32),
    "c2": dict(label="configs[2]: NeuroMechFly 18-DOF walking with leg reward, 65536 envs over 8 GPUs = 8192 per GPU, "
                     "horizon 80 / mini-chunk 5", asset="limited_18", full=True, walking=True, n=8192, horizon=80, pool=16),
    "c3": dict(label="configs[3]: full-DOF nmf_no_limits.urdf joint set (D=42, A=42, O=145), num_envs=32768 per GPU, "
                     "horizon 64", asset="full_42", full=True, walking=False, n=32768, horizon=64, pool=8),
    "c3b": dict(label="configs[3]: nmf_no_limits.urdf (D=42) with the reference's 18 actuated joints (A=18, O=73, "
                      "fly.py:21-31), num_envs=32768 per GPU, horizon 64", asset="full_42", full=False, walking=False,
                n=32768, horizon=64, pool=8),
}


def work_model(a: int, d: int):
    """Algorithmic bytes / flops per unit, as functions of A (actuated) and D (asset DOFs)."""
    o = 19 + 3 * a
    env_read = 52 + 8 * a + 11 * 12 + 4 * a + 4 + 8 + 8
    env_write = 4 * o + 4 + 8 + 8 + 4 + 4 + 24
    mlp = o * 256 + 256 * 128 + 128 * 64
    return {
        "O": o,
        "bytes_env": env_read + env_write,                                  # 764 at A=18
        "bytes_step": env_read + env_write + 4 * a + 4 + 4 * d,             # fused: + action, logp, DOF targets: 912
        "flops_act": 2 * (mlp + 64 * a),                                    # 121 600 at O=73, A=18
        "flops_value": 2 * (mlp + 64),
        "flops_update": 6 * (mlp + 64 * a + 128 * 64 + 64),                 # 414 336 per sample (trunk once)
        "bytes_gae": 20,
    }


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm": float(p["hbm_gbs"]), "hbm_src": "measured", "tf_burst": float(p["bf16_tflops"]),
                "tf_sustained": float(p["bf16_tflops_sustained"]), "tf_src": "measured"}
    return {"hbm": 6650.0, "hbm_src": "fallback", "tf_burst": 1590.0, "tf_sustained": 1400.0, "tf_src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def config_dict(cfg, world, extra=None):
    c = {"workload": cfg["label"], "num_envs_per_gpu": cfg["n"], "horizon": cfg["horizon"],
         "mini_chunk": max(1, cfg["horizon"] // 16), "env_steps_per_bench_step": cfg["n"] * cfg["horizon"] * world,
         "bench_step": "one training iteration: T-step rollout + PPO update (value forward, GAE, 5 x (T/mc - 1) "
                       "optimiser steps)",
         "parallelism": f"dp{world} (independent env shards; gradient sum in the update only)"}
    if extra:
        c.update(extra)
    return c


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_runner(cfg, threads):
    from oracle import fly_oracle as fo
    from oracle.ref_bench import ReferenceRunner
    tab = json.load(open(os.path.join(ROOT, "fly_bproject_b200", "data", "nmf_tables.json")))["assets"][cfg["asset"]]
    aidx = list(range(tab["num_dof"])) if cfg["full"] else tab["action_index_18"]
    t = fo.EnvTables(cfg["n"], tab["lower"], tab["upper"], tab["initial_pos"], aidx, tab["abdomen_local"],
                     tab["leg_tip_local"], tab["num_bodies"])
    return ReferenceRunner(t, cfg["horizon"], max(1, cfg["horizon"] // 16), cfg["walking"], threads)


def run_reference(args, cfg):
    """The reference's own CPU implementation of the path on this box's host cores (oracle/ref_bench.py): unmodified
    reference PPO.run / make_data / update + compute_fly_observations / compute_fly_reward2 when the reference's files
    travel with the snapshot (oracle/_ref), the oracle port otherwise.  Each bench step times a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    r = cpu_runner(cfg, cores)
    nsteps = max(2, min(2 * args.steps, 40))
    # per bench step: a bounded sample of the iteration = 2 env steps; the update pieces are timed once per run
    res = r.iteration(env_steps=nsteps, warmup=max(5, args.warmup), with_update=True)
    val = res["env_steps_per_s"]
    sample = (f"t_iter = T*t_env_step + t_make_data + n_mb*t_minibatch composed from: median of {nsteps} "
              f"PPO.run() calls ({res['t_env_step_s'] * 1e3:.1f} ms), one make_data over (T,N) "
              f"({res['t_make_data_s']:.2f} s), the reference update() on a 2-mini-chunk rollout = 5 optimiser steps of "
              f"{max(1, cfg['horizon'] // 16) * cfg['n']} samples ({res['t_minibatch_s']:.3f} s each); "
              f"n_mb = {res['n_minibatches']}")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": res["t_iteration_s"] * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(cfg, 1, {"note": "CPU torch, one process; env count is ONE GPU's shard"}),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": res["kind"], "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "rollout": {"metric": "fused_env_steps_per_sec", "value": res["env_steps_per_s_rollout"], "unit": UNIT},
        "update": {"metric": "ppo_update_samples_per_sec", "value": res["update_samples_per_s"], "unit": "samples/s"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ own arm
def graph_time_ms(fn, iters, replays=5):
    """Device time per launch of `fn`: `iters` launches captured into ONE CUDA graph (so the host's launch rate —
    a ctypes call costs more than some of these kernels run — is out of the picture), replayed `replays` times after
    a warm-up replay, CUDA events on the capturing stream around the replays."""
    import torch
    for i in range(min(iters, 4)):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (iters * replays)


def run_own(args, cfg):
    import torch
    import torch.distributed as dist
    import fly_bproject_b200 as fb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: flyb200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its banner / debug lines to stdout (the version line even at WARN level): send them to stderr so
        # that stdout carries the ONE JSON line and nothing else
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    if args.gpus != world and rank == 0 and world > 1:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE {world}", file=sys.stderr)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pk = peaks()
    n, horizon, pool = cfg["n"], cfg["horizon"], cfg["pool"]
    torch.manual_seed(0)
    a = fb.make_args(num_envs=n, sim_device=dev, horizon=horizon, verbose=False, state_pool=pool, seed=1234,
                     asset=cfg["asset"], full_actions=cfg["full"], walking=cfg["walking"])
    agent = fb.PPO(a)
    env = agent.env
    wm = work_model(env.num_act, env.num_dof)
    mc = agent.mini_chunk_size
    n_minibatches = agent.epoch * (agent.rollout_size // mc - 1)
    rows_mb = mc * n
    K, W, REPEATS = args.steps, max(3, args.warmup), 5
    tc_path = agent.net.wpack is not None
    fused = agent._fused_steps

    def iteration():
        agent.collect_rollout(use_graph=True)
        agent.update(use_graph=True)

    for _ in range(W):
        iteration()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()

    # ---------------- headline: K training iterations per block, REPEATS blocks, median block
    blocks = []
    for _ in range(REPEATS):
        barrier()
        e0.record()
        for _ in range(K):
            iteration()
        e1.record()
        barrier()
        blocks.append(max_over_ranks(e0.elapsed_time(e1)))
    iter_ms = statistics.median(blocks) / K
    env_steps = n * horizon * world
    value = env_steps / (iter_ms * 1e-3)

    # ---------------- the two rates the metric names, each in its own timed region (same graphs)
    def timed_blocks(fn, k):
        out = []
        for _ in range(REPEATS):
            barrier()
            e0.record()
            for _ in range(k):
                fn()
            e1.record()
            barrier()
            out.append(max_over_ranks(e0.elapsed_time(e1)) / k)
        return statistics.median(out)

    roll_ms = timed_blocks(lambda: agent.collect_rollout(use_graph=True), max(K, 10))
    upd_ms = timed_blocks(lambda: agent.update(use_graph=True), max(2, min(K, 10)))
    clocks = sampler.stop()
    reset_fraction = float(env.reset_count.item()) / n     # envs flagged by the rollout's last step (termination mix)
    roll_value = env_steps / (roll_ms * 1e-3)
    upd_value = n_minibatches * rows_mb * world / (upd_ms * 1e-3)
    step_us = roll_ms * 1e3 / horizon                              # one env step INSIDE the rollout graph
    launches_rollout = (horizon + 1) if fused else 2 * horizon
    # per optimiser step: chain, weight gradients, reduce (+ norm partials; data-parallel: + the gradient exchange),
    # clip + Adam on the tensor-core path; with the split exchange (or the CUDA-core path) a norm kernel in between
    ppo_tc = tc_path and int(agent.net.layout.k0_pad) in (80, 160)     # the two compiled shapes of the PPO step
    four = ppo_tc and (world == 1 or (agent._peers is not None and agent._dp_step == "fused"))
    # opt-in (FLYB_ADAM_FUSED=1, measured slower): clip + Adam inside the reduction when its whole grid (one CTA per 256
    # entries) can be resident at once: two CTAs of that kernel per SM (tc_ppo_grad checks the same by the occupancy API)
    adam_in = four and (int(agent.net.layout.total) + 255) // 256 <= 2 * 148 and os.environ.get("FLYB_ADAM_FUSED", "0") == "1"
    launches_update = 2 + n_minibatches * (3 if adam_in else 4 if four else 5)
    launches = K * REPEATS * (launches_rollout + launches_update)

    # ---------------- per-kernel device times, each kernel alone in a graph (pool rotation keeps L2 cold)
    import ctypes as C
    from fly_bproject_b200 import _lib
    p = _lib.ptr

    def k_env(i):
        env.simulate()
        env.post_physics(obs_copy=agent._obs_ring[(i % horizon) + 1], reward_copy=agent.all_reward[i % horizon])

    def k_act(i):
        st = env.state
        env.simulate()
        _lib.check(agent.lib.flyb_policy_act(
            C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack), p(env.obs_buf), n,
            p(agent.action_var), None, 1, i,
            None, p(agent.all_acts[i % horizon]), p(agent.all_log_prob[i % horizon]), None, None, C.byref(env.cfg),
            p(env._actions_scaled), p(env.dof_targets), p(env.reset_buf), p(env.progress_buf), p(env.potentials),
            p(env.prev_potentials), p(st.root), p(st.dof_state), _lib.current_stream()))

    def k_fused(i):
        agent._fused_step(1 + i % (horizon - 1))

    def k_gae(i):
        _lib.check(agent.lib.flyb_gae(p(agent.all_reward), p(agent._values), p(env.reset_buf), horizon, n, 0.99,
                                      0.95, p(agent._target), p(agent.all_advantage), _lib.current_stream()))

    def k_vfwd(i):
        _lib.check(agent.lib.flyb_net_forward(C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack),
                                              p(agent._obs_ring), (horizon + 1) * n, None, p(agent._values),
                                              _lib.current_stream()))

    def k_step(i):
        agent._minibatch(0, mc)

    def k_gradonly(i):
        lay = C.byref(agent.net.layout)
        _lib.check(agent.lib.flyb_ppo_grad(
            lay, C.byref(agent.ppo_cfg), p(agent.net.flat), p(agent.net.wpack), p(agent._obs_ring[0:mc]),
            p(agent.all_acts[0:mc]), p(agent.all_log_prob[0:mc]), p(agent._target[0:mc]), p(agent.all_advantage[0:mc]),
            p(agent.action_var), rows_mb, p(agent.grad), p(agent.loss), agent._ws_ptr, None, _lib.current_stream()))

    todo = [("post_physics", k_env, 4 * pool), ("policy_act", k_act, 4 * pool), ("gae_scan", k_gae, 32),
            ("value_forward", k_vfwd, 4), ("ppo_grad", k_gradonly, 8), ("ppo_minibatch_step", k_step, 8)]
    if fused:
        todo.insert(0, ("env_step_fused", k_fused, 4 * pool))
    kt = {}
    for name, fn, iters in todo:
        if world > 1 and name == "ppo_minibatch_step" and agent._peers is None:   # NCCL all-reduce: not captured
            fn(0)
            torch.cuda.synchronize()
            e0.record()
            for i in range(iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            kt[name] = e0.elapsed_time(e1) / iters
        else:
            kt[name] = graph_time_ms(fn, iters)
    where = "alone in a captured graph (pool rotation keeps the state out of L2)"
    kernels = {
        "post_physics": {"ms": kt["post_physics"], "bound": "hbm", "achieved": wm["bytes_env"] * n / kt["post_physics"] / 1e6,
                         "unit": "GB/s", "peak": pk["hbm"], "measured": where},
        "policy_act": {"ms": kt["policy_act"], "bound": "tensor", "achieved": wm["flops_act"] * n / kt["policy_act"] / 1e9,
                       "unit": "TFLOP/s", "peak": pk["tf_sustained"], "measured": where},
        "gae_scan": {"ms": kt["gae_scan"], "bound": "hbm", "achieved": wm["bytes_gae"] * n * horizon / kt["gae_scan"] / 1e6,
                     "unit": "GB/s", "peak": pk["hbm"], "measured": where},
        "value_forward": {"ms": kt["value_forward"], "bound": "tensor",
                          "achieved": wm["flops_value"] * (horizon + 1) * n / kt["value_forward"] / 1e9,
                          "unit": "TFLOP/s", "peak": pk["tf_sustained"], "measured": where},
        "ppo_grad": {"ms": kt["ppo_grad"], "bound": "tensor", "achieved": wm["flops_update"] * rows_mb / kt["ppo_grad"] / 1e9,
                     "unit": "TFLOP/s", "peak": pk["tf_sustained"], "measured": where,
                     "note": "fwd+bwd chain kernel + weight-gradient kernel + partial reduce"},
        "ppo_minibatch_step": {"ms": kt["ppo_minibatch_step"], "bound": "tensor",
                               "achieved": wm["flops_update"] * rows_mb / kt["ppo_minibatch_step"] / 1e9,
                               "unit": "TFLOP/s", "peak": pk["tf_sustained"], "measured": where,
                               "note": "ppo_grad + clip/Adam + in-place weight-stream refresh (+ gradient exchange at N > 1)"},
    }
    # the env step as it runs in the HEADLINE's rollout graph: rollout time / T (launch gaps included)
    kernels["env_step_in_rollout"] = {
        "ms": step_us * 1e-3, "bound": "tensor", "achieved": wm["flops_act"] * n / step_us / 1e6,
        "unit": "TFLOP/s", "peak": pk["tf_sustained"],
        "measured": "inside the timed rollout graph: rollout time / T",
        "launches_per_env_step": launches_rollout / horizon,
        "hbm_gbs": wm["bytes_step"] * n / step_us / 1e3, "hbm_frac": wm["bytes_step"] * n / step_us / 1e3 / pk["hbm"]}
    if fused:
        kernels["env_step_fused"] = {"ms": kt["env_step_fused"], "bound": "tensor",
                                     "achieved": wm["flops_act"] * n / kt["env_step_fused"] / 1e9, "unit": "TFLOP/s",
                                     "peak": pk["tf_sustained"], "measured": where}
    for k in kernels.values():
        k["frac"] = k["achieved"] / k["peak"]
        if k["bound"] == "tensor":
            if tc_path:   # 3xTF32: TF32 runs at half the bf16 rate and every product takes 3 MMAs
                k["mma_kind"] = MMA_KIND
                k["kind_peak"] = k["peak"] / 6.0
                k["frac_of_kind_peak"] = k["achieved"] / k["kind_peak"]
            else:
                k["mma_kind"] = "none: CUDA-core fp32 FFMA path (shape outside the tensor-core layout)"
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    # Dominant work of the bench step (the training iteration): the optimiser step — 5 x (T/mc - 1) of them are ~90 % of
    # the iteration, and within one the forward/backward chain kernel and the weight-gradient kernel are ~50 % / ~44 %
    # of the kernel time (ncu launch list under profiles/).  Timed live: CUDA events over a captured graph of 8 steps.
    roofline = dict(kernels["ppo_minibatch_step"])
    roofline.update({
        "kernel": ("ppo_fwd_bwd_tc_kernel + ppo_dw_tc_kernel (+ ppo_reduce_tc_kernel, ppo_adam_kernel): one optimiser step"
                   if tc_path else "ppo_fwd_bwd_kernel + ppo_dw_kernel (+ reduce, norm, Adam): one optimiser step"),
        "traffic": traffic.get("ppo_minibatch_step"),
        "traffic_source": ("profiles/traffic.json <- " + str(traffic.get("_source", "profiles/r03_hot_metrics.txt")) +
                           " (one ncu --set full capture, not measured by this run)") if traffic else None,
        "algorithmic_per_launch": {"flops": wm["flops_update"] * rows_mb, "samples": rows_mb},
        "share_of_bench_step": n_minibatches * kt["ppo_minibatch_step"] / iter_ms,
        "peak_source": pk["tf_src"] + " (bf16 sustained: timed inside a long step)"})
    roofline.pop("ms", None)
    roofline.pop("note", None)
    # the rollout's dominant kernel: the (fused) env step, timed live inside the rollout graph
    dom = "env_step_fused" if fused else ("policy_act" if kt["policy_act"] >= kt["post_physics"] else "post_physics")
    roofline_rollout = dict(kernels["env_step_in_rollout"] if fused else kernels[dom])
    roofline_rollout.update({"kernel": "env_step_tc_kernel" if fused else dom, "traffic": traffic.get(dom),
                             "algorithmic_per_launch": {"flops": wm["flops_act"] * n, "bytes": wm["bytes_step"] * n}})
    roofline_rollout.pop("ms", None)

    # ---------------- e2e: public API with HOST state buffers, H2D + D2H inside the timed region
    n_host = 4
    host_pool = [{k: v.cpu().pin_memory() for k, v in e.items()} for e in env.state.pool[:n_host]]
    live = {k: torch.empty_like(v) for k, v in env.state.pool[0].items()}
    step_no = [0]

    # The caller's loop is double-buffered over the two PCIe directions: the DOF targets, reset mask, ids and count (what a
    # host-resident simulator needs before it can advance) are read back on the compute stream and waited for; the
    # observation / reward read-back (4.8 MB of the 6.2) rides a second stream UNDER the next step's H2D state copy, and
    # the kernels that overwrite obs_buf / reward_buf wait for it (wait_event at the end of h2d).  Every byte still moves
    # inside the timed region, every step.
    side = torch.cuda.Stream()
    ev_step, ev_side = torch.cuda.Event(), torch.cuda.Event()
    ev_side.record(side)

    def h2d(src):
        hp = host_pool[step_no[0] % n_host]
        for k in ("root", "dof_state", "force"):
            live[k].copy_(hp[k], non_blocking=True)
        torch.cuda.current_stream().wait_event(ev_side)    # post_physics may not overwrite obs / reward before they left
        step_no[0] += 1

    a2 = fb.make_args(num_envs=n, sim_device=dev, horizon=horizon, verbose=False, seed=99, asset=cfg["asset"],
                      full_actions=cfg["full"], walking=cfg["walking"])
    src = fb.TensorStateSource(live["root"], live["dof_state"], live["force"], simulate_fn=h2d)
    agent2 = fb.PPO(a2, env=fb.Fly(a2, state_source=src))
    agent2.net.load_state_dict(agent.net.state_dict())
    env2 = agent2.env
    back = {"obs": env2.obs_buf, "reward": env2.reward_buf, "reset": env2.reset_buf, "reset_ids": env2.reset_ids,
            "reset_count": env2.reset_count, "dof_targets": env2.dof_targets}
    late = ("obs", "reward")                               # written by post_physics only (fly.py:676-681)
    host_back = [{k: torch.empty(v.shape, dtype=v.dtype).pin_memory() for k, v in back.items()} for _ in range(2)]

    def e2e_step():
        hb = host_back[step_no[0] & 1]
        agent2.run()                                       # the T-th call of a rollout also runs the update
        ev_side.synchronize()                              # the PREVIOUS step's obs / reward are on the host by now
        main = torch.cuda.current_stream()
        ev_step.record(main)
        side.wait_event(ev_step)
        with torch.cuda.stream(side):
            for k in late:
                hb[k].copy_(back[k], non_blocking=True)
            ev_side.record(side)
        for k, v in back.items():
            if k not in late:
                hb[k].copy_(v, non_blocking=True)
        main.synchronize()                                 # the simulator consumes targets / reset ids before it advances

    for _ in range(horizon):                               # one warm-up iteration (includes an update)
        e2e_step()
    ke = max(1, min(2, 256 // horizon)) * horizon          # whole iterations
    barrier()
    e0.record()
    for _ in range(ke):
        e2e_step()
    torch.cuda.current_stream().wait_event(ev_side)        # the last read-back ends inside the timed region
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    h2d_bytes = sum(v.numel() * v.element_size() for v in live.values())
    d2h_bytes = sum(v.numel() * v.element_size() for v in back.values())
    e2e_value = n * ke * world / (e2e_ms * 1e-3)

    # N > 1: a correctness bit for the data-parallel update.  After the timed iterations every rank must hold the same
    # parameters and Adam state (the gradient sum is the only thing the ranks share) and no exchange may have timed out.
    dp_check = None
    if world > 1:
        torch.cuda.synchronize()

        def digest(t):      # position-weighted sum of the raw bits, mod 2^64
            bits = t.contiguous().view(torch.int32).to(torch.int64)
            return (bits * torch.arange(1, bits.numel() + 1, device=bits.device, dtype=torch.int64)).sum()

        status = agent._peers.status() if agent._peers is not None else 0
        mine = torch.stack([digest(agent.net.flat), digest(agent.exp_avg), digest(agent.exp_avg_sq),
                            torch.tensor(int(status), dtype=torch.int64, device=dev),
                            torch.tensor(int(agent.optim_step), dtype=torch.int64, device=dev)])
        everyone = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(everyone, mine)
        everyone = torch.stack(everyone).cpu()
        dp_check = {"identical": bool((everyone[:, [0, 1, 2, 4]] == everyone[0, [0, 1, 2, 4]]).all()),
                    "peer_status": int(everyone[:, 3].max()), "optimiser_steps": int(everyone[0, 4]),
                    "finite": bool(torch.isfinite(agent.net.flat).all()),
                    "what": "64-bit checksums of parameters, exp_avg, exp_avg_sq and the step count all-gathered after "
                            "the timed region: equal on every rank; sticky peer-exchange error word 0 on every rank"}

    if rank == 0:
        if env.state.pool_bytes > 126e6:
            l2 = (f"inputs larger than L2: pool of {pool} synthetic states = {env.state.pool_bytes / 1e6:.0f} MB, rotated "
                  "every env-step")
        else:
            l2 = (f"state pool of {pool} entries = {env.state.pool_bytes / 1e6:.0f} MB is L2-resident at this env count "
                  "(reported, not flushed: the reference's own small case)")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": iter_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_dict(cfg, world, {
                "l2": l2, "rng": "in-kernel Philox4x32-10",
                "gemm_path": MMA_KIND if tc_path else "CUDA-core fp32 (shape outside the tensor-core layout)",
                "timing": f"median of {REPEATS} blocks of {K} iterations, CUDA events, max over ranks",
                "block_ms": blocks}),
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "env_steps": ke, "ms_per_env_step": e2e_ms / ke, "h2d_gbs_per_rank": h2d_bytes * ke / e2e_ms / 1e6,
                    "api": "PPO.run() per env-step (the T-th call runs the update): state tensors H2D from pinned host "
                           "memory; DOF targets, reset mask, compacted ids + count read back D2H and waited for, obs + reward read back "
                           "on a second stream under the next step's H2D (double-buffered pinned buffers)"},
            "roofline": roofline, "roofline_rollout": roofline_rollout, "kernels": kernels,
            "rollout": {"metric": "fused_env_steps_per_sec", "value": roll_value, "unit": UNIT, "ms_per_rollout": roll_ms,
                        "us_per_env_step": step_us, "launches_per_rollout": launches_rollout,
                        "reset_fraction_last_step": reset_fraction,
                        "includes": ("obs + reward + reset + id compaction + policy act + action scaling + reset "
                                     "bookkeeping, ") + ("one launch per env step" if fused else "two launches per env step")},
            "update": {"metric": "ppo_update_samples_per_sec", "value": upd_value, "unit": "samples/s",
                       "ms_per_update": upd_ms, "optimiser_steps": n_minibatches, "samples_per_step": rows_mb,
                       "us_per_optimiser_step_isolated": kt["ppo_minibatch_step"] * 1e3,
                       "includes": "value forward over (T+1)N rows + GAE scan + optimiser steps (fwd/bwd/clip/Adam)" +
                                   ("" if world == 1 else " + NCCL all-reduce" if agent._peers is None else
                                    " with the gradient sum over NVLink peer memory inside the gradient-reducing kernel (FLYB_DP_STEP=" +
                                    agent._dp_step + ")")},
        }
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            res = cpu_runner(cfg, cores).iteration(env_steps=12, warmup=4, with_update=True)
            line["cpu_baseline"] = {
                "value": res["env_steps_per_s"], "unit": UNIT, "cores": cores, "kind": res["kind"],
                "rollout_env_steps_per_s": res["env_steps_per_s_rollout"],
                "update_samples_per_s": res["update_samples_per_s"],
                "sample": f"iteration composed from bounded samples (oracle/ref_bench.py): 12 PPO.run() calls "
                          f"({res['t_env_step_s'] * 1e3:.1f} ms median), one make_data ({res['t_make_data_s']:.2f} s), "
                          f"5 optimiser steps of {rows_mb} samples ({res['t_minibatch_s']:.3f} s each) x {n_minibatches}"}
        if dp_check is not None:
            line["dp_check"] = dp_check
        print(json.dumps(line), flush=True)
    if world > 1:
        agent.exit()            # unmaps the peers' exchange buffers (collective)
        dist.destroy_process_group()
        if not (dp_check["identical"] and dp_check["peer_status"] == 0 and dp_check["finite"]):
            print(f"bench.py: data-parallel check FAILED on rank {rank}: {dp_check}", file=sys.stderr)
            sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", type=str, default="flyb200", choices=["flyb200", "reference"])
    ap.add_argument("--config", type=str, default="c1", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu", dest="no_cpu", action="store_true", help="skip the cpu_baseline leg (sweeps)")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg)
    else:
        run_own(args, cfg)


if __name__ == "__main__":
    main()
