#!/usr/bin/env python
"""Benchmark of the fly_bProject hot path on B200 (contract: see the task brief / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload = BASELINE.json configs[1]: NeuroMechFly 18-DOF standing, num_envs = 16384 per GPU, horizon 32.
One bench STEP = one rollout pass of the hot path over the env batch: T = 32 fused env-steps, each
[policy act + action scaling + reset bookkeeping] -> [synthetic Isaac Gym state] -> [obs + reward + reset mask + id
compaction], replayed as one CUDA graph.  `value` = env-steps/s, whole job, inputs resident in HBM (a pool of 16
synthetic states = 344 MB > L2, rotated every env-step).  The PPO update (GAE + 75 optimiser steps) is timed in its
own region and reported under "update" (samples/s) — BASELINE.json's metric names both rates.
`e2e` = the same env-steps/s through the public API (`PPO.run()`), per env-step: H2D copy of that step's state tensors
from pinned host memory, the kernels, D2H read of the step's reward / reset count.
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

NUM_ENVS = 16384
HORIZON = 32
POOL = 16
METRIC = "env_steps_per_sec"
UNIT = "env-steps/s"
WORKLOAD = "configs[1]: NeuroMechFly 18-DOF standing (A=D=18, O=73), num_envs=16384 per GPU, horizon 32"

# algorithmic bytes / flops per unit
A, D, O, B = 18, 18, 73, 93
BYTES_ENV_KERNEL = (52 + 8 * A + 11 * 12 + 4 * A + 4 + 8 + 8) + (4 * O + 4 + 8 + 8 + 4 + 4 + 24)          # 764
FLOPS_ACT = 2 * (O * 256 + 256 * 128 + 128 * 64 + 64 * A)                                              # 121 600
BYTES_ACT = 4 * O + 4 * A + 4 + 4 * D + 4 * A                                                           # obs read + action/logp/targets/scaled
FLOPS_UPDATE = 6 * (O * 256 + 256 * 128 + 128 * 64 + 64 * A + 128 * 64 + 64)                            # 414 336 per sample
BYTES_GAE = 20
MMA_KIND = "tcgen05 kind::tf32, 3xTF32 (hi*lo + lo*hi + hi*hi), fp32 accumulate in TMEM"


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


# ------------------------------------------------------------------------------------------------ reference arm
def oracle_env_step_setup(n, threads):
    import torch
    from oracle import fly_oracle as fo
    torch.set_num_threads(threads)
    tab = json.load(open(os.path.join(ROOT, "fly_bproject_b200", "data", "nmf_tables.json")))["assets"]["limited_18"]
    t = fo.EnvTables(n, tab["lower"], tab["upper"], tab["initial_pos"], list(range(18)), tab["abdomen_local"],
                     tab["leg_tip_local"], tab["num_bodies"])
    states = [fo.synth_state(t, 1234 + 7919 * i) for i in range(2)]
    s = fo.EnvState(t)
    s.reset_buf.zero_()
    p = fo.init_params(73, 18, seed=0)
    var = torch.full((18,), 0.2)
    gen = torch.Generator().manual_seed(0)

    def step(i):
        st = states[i % 2]
        noise = torch.randn(n, 18, generator=gen)
        a, logp, _ = fo.act(p, s.obs_buf, var, noise)                 # ppo.py:213-220
        s.actions, _ = fo.scale_actions(t, a)                          # fly.py:626-657
        fo.apply_reset(t, s)                                           # fly.py:446-480
        s.root, s.dof_states, s.force = st["root"], st["dof_states"], st["force"]
        fo.post_physics(t, s)                                          # fly.py:676-681
    return step


def time_oracle_env_steps(n, steps, warmup, threads):
    step = oracle_env_step_setup(n, threads)
    for i in range(warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    dt = time.perf_counter() - t0
    return n * steps / dt, dt / steps


def time_oracle_update(n, mc, threads, minibatches=2):
    import torch
    from oracle import fly_oracle as fo
    torch.set_num_threads(threads)
    gen = torch.Generator().manual_seed(0)
    p = fo.init_params(73, 18, seed=0)
    st = fo.AdamState(p)
    rows = mc * n
    obs = torch.randn(mc, n, 73, generator=gen)
    act = torch.randn(mc, n, 18, generator=gen).clip(-1, 1)
    logp = torch.randn(mc, n, generator=gen) - 20
    target = torch.randn(mc, n, 1, generator=gen)
    adv = torch.randn(mc, n, 1, generator=gen)
    var = torch.full((18,), 0.2)
    fo.minibatch_step(p, st, obs, act, logp, target, adv, var)
    t0 = time.perf_counter()
    for _ in range(minibatches):
        fo.minibatch_step(p, st, obs, act, logp, target, adv, var)
    dt = time.perf_counter() - t0
    return rows * minibatches / dt


def run_reference(args):
    """The reference's own CPU implementation of the path, timed on this box's host cores (oracle port of the
    reference's torch code — the reference checkout itself does not exist on the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    threads = max(1, cores)
    # one bench step = T env-steps in our arm; the CPU arm times a bounded sample: `steps` x 1 env-step of the batch
    val, per = time_oracle_env_steps(NUM_ENVS, max(1, args.steps), max(1, args.warmup), threads)
    upd = time_oracle_update(NUM_ENVS, 2, threads, minibatches=1)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU torch oracle port of reference fly.py/ppo.py hot path; one step = "
                   "one env-step of the 16384-env batch (act + scale + reset + obs + reward)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{args.steps} env-steps x {NUM_ENVS} envs after {args.warmup} warm-up"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "update": {"metric": "ppo_update_samples_per_sec", "value": upd, "unit": "samples/s",
                   "sample": "1 optimiser step x 32768 samples"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ own arm
def cuda_time_ms(fn, iters, stream=None):
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


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


def run_own(args):
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
        # (NCCL honours NCCL_DEBUG_FILE only above the VERSION level, so VERSION is raised to WARN — same banner)
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
    torch.manual_seed(0)
    a = fb.make_args(num_envs=NUM_ENVS, sim_device=dev, horizon=HORIZON, verbose=False, state_pool=POOL,
                     seed=1234)
    agent = fb.PPO(a)
    env = agent.env
    K, W = args.steps, max(3, args.warmup)

    # ---------------- rollout (headline): K graph replays of T env-steps
    for _ in range(W):
        agent.collect_rollout(use_graph=True)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(K):
        agent.collect_rollout(use_graph=True)
    e1.record()
    barrier()
    roll_ms = max_over_ranks(e0.elapsed_time(e1))
    env_steps = NUM_ENVS * HORIZON * K * world
    value = env_steps / (roll_ms * 1e-3)
    launches = K * HORIZON * 2

    # ---------------- update: GAE + 75 optimiser steps (graph), timed on its own
    for _ in range(2):
        agent.update(use_graph=True)
    ku = max(2, min(K, 10))
    barrier()
    e0.record()
    for _ in range(ku):
        agent.update(use_graph=True)
    e1.record()
    barrier()
    upd_ms = max_over_ranks(e0.elapsed_time(e1)) / ku
    clocks = sampler.stop()
    n_minibatches = agent.epoch * (agent.rollout_size // agent.mini_chunk_size - 1)
    upd_samples = n_minibatches * agent.mini_chunk_size * NUM_ENVS * world
    upd_value = upd_samples / (upd_ms * 1e-3)

    # ---------------- per-kernel device times (CUDA events on the launching stream; pool rotation keeps L2 cold)
    import ctypes as C
    from fly_bproject_b200 import _lib
    p = _lib.ptr

    def k_env(i):
        env.simulate()
        env.post_physics(obs_copy=agent._obs_ring[(i % HORIZON) + 1], reward_copy=agent.all_reward[i % HORIZON])

    def k_act(i):
        st = env.state
        env.simulate()
        _lib.check(agent.lib.flyb_policy_act(
            C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack), p(env.obs_buf), NUM_ENVS,
            p(agent.action_var), None, 1, i,
            None, p(agent.all_acts[i % HORIZON]), p(agent.all_log_prob[i % HORIZON]), None, None, C.byref(env.cfg),
            p(env._actions_scaled), p(env.dof_targets), p(env.reset_buf), p(env.progress_buf), p(env.potentials),
            p(env.prev_potentials), p(st.root), p(st.dof_state), _lib.current_stream()))

    def k_gae(i):
        _lib.check(agent.lib.flyb_gae(p(agent.all_reward), p(agent._values), p(env.reset_buf), HORIZON, NUM_ENVS, 0.99,
                                      0.95, p(agent._target), p(agent.all_advantage), _lib.current_stream()))

    def k_vfwd(i):
        _lib.check(agent.lib.flyb_net_forward(C.byref(agent.net.layout), p(agent.net.flat), p(agent.net.wpack),
                                              p(agent._obs_ring),
                                              (HORIZON + 1) * NUM_ENVS, None, p(agent._values), _lib.current_stream()))

    def k_grad(i):
        agent._minibatch(0, agent.mini_chunk_size)

    def k_gradonly(i):
        lay = C.byref(agent.net.layout)
        _lib.check(agent.lib.flyb_ppo_grad(
            lay, C.byref(agent.ppo_cfg), p(agent.net.flat), p(agent.net.wpack), p(agent._obs_ring[0:agent.mini_chunk_size]),
            p(agent.all_acts[0:agent.mini_chunk_size]), p(agent.all_log_prob[0:agent.mini_chunk_size]),
            p(agent._target[0:agent.mini_chunk_size]), p(agent.all_advantage[0:agent.mini_chunk_size]),
            p(agent.action_var), agent.mini_chunk_size * NUM_ENVS, p(agent.grad), p(agent.loss), agent._ws_ptr,
            None, _lib.current_stream()))

    kt = {}
    for name, fn, iters in [("post_physics", k_env, 4 * POOL), ("policy_act", k_act, 4 * POOL), ("gae_scan", k_gae, 32),
                            ("value_forward", k_vfwd, 4), ("ppo_grad", k_gradonly, 8),
                            ("ppo_minibatch_step", k_grad, 8)]:
        if world > 1 and name == "ppo_minibatch_step" and agent._peers is None:   # NCCL all-reduce: not captured
            fn(0)
            kt[name] = cuda_time_ms(fn, iters)
        else:
            kt[name] = graph_time_ms(fn, iters)
    rows_mb = agent.mini_chunk_size * NUM_ENVS
    kernels = {
        "post_physics": {"ms": kt["post_physics"], "bound": "hbm",
                         "achieved": BYTES_ENV_KERNEL * NUM_ENVS / kt["post_physics"] / 1e6, "unit": "GB/s",
                         "peak": pk["hbm"]},
        "policy_act": {"ms": kt["policy_act"], "bound": "tensor",
                       "achieved": FLOPS_ACT * NUM_ENVS / kt["policy_act"] / 1e9, "unit": "TFLOP/s",
                       "peak": pk["tf_sustained"], "mma_kind": MMA_KIND},
        "gae_scan": {"ms": kt["gae_scan"], "bound": "hbm",
                     "achieved": BYTES_GAE * NUM_ENVS * HORIZON / kt["gae_scan"] / 1e6, "unit": "GB/s", "peak": pk["hbm"]},
        "value_forward": {"ms": kt["value_forward"], "bound": "tensor",
                          "achieved": 2 * (O * 256 + 256 * 128 + 128 * 64 + 64) * (HORIZON + 1) * NUM_ENVS /
                          kt["value_forward"] / 1e9, "unit": "TFLOP/s", "peak": pk["tf_sustained"], "mma_kind": MMA_KIND},
        "ppo_grad": {"ms": kt["ppo_grad"], "bound": "tensor",
                     "achieved": FLOPS_UPDATE * rows_mb / kt["ppo_grad"] / 1e9, "unit": "TFLOP/s",
                     "peak": pk["tf_sustained"], "mma_kind": MMA_KIND,
                     "note": "fwd+bwd chain kernel + weight-gradient kernel + partial reduce"},
        "ppo_minibatch_step": {"ms": kt["ppo_minibatch_step"], "bound": "tensor",
                               "achieved": FLOPS_UPDATE * rows_mb / kt["ppo_minibatch_step"] / 1e9, "unit": "TFLOP/s",
                               "peak": pk["tf_sustained"], "mma_kind": MMA_KIND,
                               "note": "ppo_grad + norm + clip/Adam + weight re-pack"},
    }
    for k in kernels.values():
        k["frac"] = k["achieved"] / k["peak"]
        if k["bound"] == "tensor":   # 3xTF32: TF32 runs at half the bf16 rate and every product takes 3 MMAs
            k["kind_peak"] = k["peak"] / 6.0
            k["frac_of_kind_peak"] = k["achieved"] / k["kind_peak"]
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    dom = "policy_act" if kt["policy_act"] >= kt["post_physics"] else "post_physics"
    roofline = dict(kernels[dom])
    roofline.update({"kernel": dom, "traffic": traffic.get(dom),
                     "peak_source": pk["tf_src" if roofline["bound"] == "tensor" else "hbm_src"] +
                     (" (sustained: timed inside a long step)" if roofline["bound"] == "tensor" else "")})
    roofline.pop("ms", None)

    # ---------------- e2e: public API with HOST state buffers, H2D + D2H inside the timed region
    n_host = 4
    host_pool = [{k: v.cpu().pin_memory() for k, v in e.items()} for e in env.state.pool[:n_host]]
    live = {k: torch.empty_like(v) for k, v in env.state.pool[0].items()}
    step_no = [0]

    def h2d(src):
        hp = host_pool[step_no[0] % n_host]
        for k in ("root", "dof_state", "force"):
            live[k].copy_(hp[k], non_blocking=True)
        step_no[0] += 1

    a2 = fb.make_args(num_envs=NUM_ENVS, sim_device=dev, horizon=HORIZON, verbose=False, testing=True, seed=99)
    src = fb.TensorStateSource(live["root"], live["dof_state"], live["force"], simulate_fn=h2d)
    agent2 = fb.PPO(a2, env=fb.Fly(a2, state_source=src))
    agent2.net.load_state_dict(agent.net.state_dict())
    host_reward = torch.empty(NUM_ENVS, dtype=torch.float32).pin_memory()
    host_count = torch.empty(1, dtype=torch.int32).pin_memory()

    def e2e_step(i):
        agent2.run()
        host_reward.copy_(agent2.env.reward_buf, non_blocking=True)
        host_count.copy_(agent2.env.reset_count, non_blocking=True)
        torch.cuda.current_stream().synchronize()          # the caller consumes the step's result

    for i in range(5):
        e2e_step(i)
    ke = 4 * HORIZON
    barrier()
    e2e_ms = max_over_ranks(cuda_time_ms(e2e_step, ke) * ke)   # eager on purpose: the user-facing call, host in the loop
    h2d_bytes = sum(v.numel() * v.element_size() for v in live.values())
    d2h_bytes = host_reward.numel() * 4 + 4
    e2e_value = NUM_ENVS * ke * world / (e2e_ms * 1e-3)

    if rank == 0:
        cores = os.cpu_count() or 1
        cpu_val, cpu_per = time_oracle_env_steps(NUM_ENVS, 40, 5, cores) if world == 1 else (None, None)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": roll_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "num_envs_per_gpu": NUM_ENVS, "horizon": HORIZON,
                       "env_steps_per_bench_step": NUM_ENVS * HORIZON,
                       "l2": f"inputs larger than L2: pool of {POOL} synthetic states = {env.state.pool_bytes / 1e6:.0f} MB, "
                             "rotated every env-step", "parallelism": f"dp{world} (independent env shards; gradient "
                             "sum in the update only)", "rng": "in-kernel Philox4x32-10"},
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "steps": ke, "api": "PPO.run() per env-step, state tensors from pinned host memory"},
            "roofline": roofline, "kernels": kernels,
            "update": {"metric": "ppo_update_samples_per_sec", "value": upd_value, "unit": "samples/s",
                       "ms_per_update": upd_ms, "optimiser_steps": n_minibatches, "samples_per_step": rows_mb,
                       "includes": "value forward over (T+1)N rows + GAE scan + 75 x (fwd/bwd/clip/Adam)" +
                                   ("" if world == 1 else " + NCCL all-reduce" if agent._peers is None else
                                    " with the gradient sum over NVLink peer memory inside the reduce/norm kernels")},
        }
        if cpu_val is not None:
            line["cpu_baseline"] = {"value": cpu_val, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"40 env-steps x {NUM_ENVS} envs (oracle: act + scale + reset + obs + reward), "
                                              f"{cpu_per * 1e3:.1f} ms/step"}
        print(json.dumps(line), flush=True)
    if world > 1:
        agent.exit()            # unmaps the peers' exchange buffers (collective)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", type=str, default="flyb200", choices=["flyb200", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
