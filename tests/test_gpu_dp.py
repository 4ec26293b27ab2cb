"""Data-parallel PPO update on real GPUs (NCCL): needs >= 2 CUDA devices, skipped otherwise.

Two ranks, each with its own env shard and RNG stream but the same initial parameters, run one rollout + two full
updates (150 optimiser steps, gradient sum inside flyb_ppo_step_dp, or between flyb_ppo_grad and flyb_ppo_apply).  Every rank must end with
bit-identical parameters / Adam state.  The sum runs once through the library's own peer-memory kernels (captured in a
CUDA graph and replayed) and once through NCCL (eager): with two ranks a + b is one rounding either way, so the two
exchanges must agree bit for bit.
"""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    """mp.spawn with a fresh rendezvous port; the port probe can lose a race against another socket on the box."""
    for attempt in range(3):
        try:
            mp.spawn(fn, args=(world, _free_port()) + args, nprocs=world, join=True)
            return
        except Exception as e:      # noqa: BLE001
            if "address already in use" not in str(e).lower() or attempt == 2:
                raise


def _worker(rank, world, port, out, mode):
    import torch.distributed as dist
    import fly_bproject_b200 as fb
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    wide = mode.startswith("wide_")          # the 42-joint set: the wide tensor-core shape's kernels (O = 145, A = 42)
    mode = mode[5:] if wide else mode
    os.environ["FLYB_DP"] = "nccl" if mode == "nccl" else "peer"
    os.environ["FLYB_DP_TAIL"] = "1" if mode == "peer_tail" else "0"
    os.environ["FLYB_DP_STEP"] = "split" if mode == "peer_split" else "fused"
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    try:
        torch.manual_seed(100 + rank)          # different initial nets per rank: the broadcast must align them
        args = fb.make_args(num_envs=2048, sim_device=f"cuda:{rank}", horizon=16, verbose=False, state_pool=4,
                            seed=1234,           # IDENTICAL args on every rank: the library decorrelates the shards
                            **({"asset": "full_42", "full_actions": True} if wide else {}))
        agent = fb.PPO(args)
        assert agent.net.wpack is not None and agent.net.num_act == (42 if wide else 18)
        assert agent.world_size == world and (agent._peers is not None) == (mode != "nccl")
        assert agent._dp_step == ("split" if mode == "peer_split" else "fused")
        agent.collect_rollout(use_graph=False)
        acts_sum, obs_sum = float(agent.all_acts.double().sum()), float(agent._obs_ring.double().sum())
        for _ in range(2):      # peer modes: eager warm-up + capture, then a replay; NCCL mode: eager twice
            agent.update(use_graph=(mode != "nccl"))
        torch.cuda.synchronize()
        torch.save({"flat": agent.net.flat.cpu(), "m": agent.exp_avg.cpu(), "v": agent.exp_avg_sq.cpu(),
                    "steps": agent.optim_step, "loss": float(agent.loss), "acts_sum": acts_sum, "obs_sum": obs_sum,
                    "peer_status": agent._peers.status() if agent._peers is not None else 0},
                   f"{out}.{'wide_' if wide else ''}{mode}.{rank}")
        agent.exit()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_data_parallel_update_keeps_ranks_identical(tmp_path):
    world = int(os.environ.get("FLYB_TEST_WORLD", "2"))       # 4 / 8 for a wider box; more ranks, same checks
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    out = str(tmp_path / "dp")
    res = {}
    # peer: exchange inside the reducing kernel, self-validating words (default); peer_split: push + fence, then flags +
    # sum in the norm kernel; peer_tail: the exchange inside the weight-gradient kernel's cooperative tail
    for mode in ("peer", "peer_split", "nccl", "peer_tail", "wide_peer", "wide_peer_split"):
        _spawn(_worker, world, out, mode)
        rs = [torch.load(f"{out}.{mode}.{r}") for r in range(world)]
        for r in rs[1:]:
            assert r["steps"] == rs[0]["steps"] == 150
            for k in ("flat", "m", "v"):
                assert torch.equal(r[k], rs[0][k]), (mode, k)
            assert r["loss"] != rs[0]["loss"]          # each rank logs its own shard's loss
            # same args on every rank, different state pool and action noise (dp.rank_seed): no duplicated shards
            assert r["acts_sum"] != rs[0]["acts_sum"] and r["obs_sum"] != rs[0]["obs_sum"]
        assert all(r["peer_status"] == 0 for r in rs)
        assert bool(torch.isfinite(rs[0]["flat"]).all())
        res[mode] = rs[0]
    for k in ("flat", "m", "v"):
        if world == 2:      # a + b is one rounding in any order (and both run the same norm kernel)
            assert torch.equal(res["peer_split"][k], res["nccl"][k]), k
    # more ranks: NCCL's tree / ring order differs from the rank order, and 150 Adam steps (each ~lr in size whatever
    # the gradient's magnitude) amplify the reassociation noise: only the bulk has to agree
    assert float((res["peer_split"]["flat"] - res["nccl"]["flat"]).abs().mean()) < 1e-3
    # same sums in the same order; only the association of the norm's partial sums differs (an ulp in the clip coefficient)
    assert float((res["peer_tail"]["flat"] - res["peer_split"]["flat"]).abs().mean()) < 1e-3
    assert float((res["peer"]["flat"] - res["peer_split"]["flat"]).abs().mean()) < 1e-3
    assert float((res["wide_peer"]["flat"] - res["wide_peer_split"]["flat"]).abs().mean()) < 1e-3


def _soak_worker(rank, world, port, out, exchanges):
    import torch.distributed as dist
    import fly_bproject_b200 as fb
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ["FLYB_DP"] = "peer"
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    try:
        torch.manual_seed(7)
        args = fb.make_args(num_envs=1024, sim_device=f"cuda:{rank}", horizon=16, verbose=False, state_pool=2, seed=1)
        agent = fb.PPO(args)
        assert agent._peers is not None
        agent.collect_rollout(use_graph=False)
        updates = (exchanges + 74) // 75
        for u in range(updates):          # one captured graph of 75 exchanges, replayed; ranks drift freely in between
            agent.update(use_graph=True)
            if u % 16 == rank:            # deliberate skew: one rank at a time stalls its host (and so its stream)
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        st = agent._peers.status()
        torch.save({"flat": agent.net.flat.cpu(), "m": agent.exp_avg.cpu(), "v": agent.exp_avg_sq.cpu(),
                    "steps": agent.optim_step, "status": st}, f"{out}.soak.{rank}")
        agent.exit()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_peer_exchange_soak(tmp_path):
    """>= 10 000 gradient exchanges over peer memory (FLYB_SOAK_EXCHANGES), every rank bit-identical at the end, no
    exchange timed out (sticky status word 0)."""
    world = int(os.environ.get("FLYB_TEST_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    exchanges = int(os.environ.get("FLYB_SOAK_EXCHANGES", "10000"))
    out = str(tmp_path / "dp")
    _spawn(_soak_worker, world, out, exchanges)
    rs = [torch.load(f"{out}.soak.{r}") for r in range(world)]
    assert rs[0]["steps"] >= exchanges
    for r in rs:
        assert r["status"] == 0
        assert r["steps"] == rs[0]["steps"]
        for k in ("flat", "m", "v"):
            assert torch.equal(r[k], rs[0][k]), k
    assert bool(torch.isfinite(rs[0]["flat"]).all())
    print(f"soak: {rs[0]['steps']} exchanges x {world} ranks, parameters / Adam state bit-identical on every rank")
