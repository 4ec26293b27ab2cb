"""Data-parallel PPO update on real GPUs (NCCL): needs >= 2 CUDA devices, skipped otherwise.

Two ranks, each with its own env shard and RNG stream but the same initial parameters, run one rollout + two full
updates (150 optimiser steps, gradient sum between flyb_ppo_grad and flyb_ppo_apply).  Every rank must end with
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


def _worker(rank, world, port, out, mode):
    import torch.distributed as dist
    import fly_bproject_b200 as fb
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ["FLYB_DP"] = mode
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    try:
        torch.manual_seed(100 + rank)          # different initial nets per rank: the broadcast must align them
        args = fb.make_args(num_envs=2048, sim_device=f"cuda:{rank}", horizon=16, verbose=False, state_pool=4,
                            seed=1234 + rank)
        agent = fb.PPO(args)
        assert agent.world_size == world and (agent._peers is not None) == (mode == "peer")
        agent.collect_rollout(use_graph=False)
        for _ in range(2):      # peer mode: eager warm-up + capture, then a replay; NCCL mode: eager twice
            agent.update(use_graph=(mode == "peer"))
        torch.cuda.synchronize()
        torch.save({"flat": agent.net.flat.cpu(), "m": agent.exp_avg.cpu(), "v": agent.exp_avg_sq.cpu(),
                    "steps": agent.optim_step, "loss": float(agent.loss)}, f"{out}.{mode}.{rank}")
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
    for mode in ("peer", "nccl"):
        mp.spawn(_worker, args=(world, _free_port(), out, mode), nprocs=world, join=True)
        rs = [torch.load(f"{out}.{mode}.{r}") for r in range(world)]
        for r in rs[1:]:
            assert r["steps"] == rs[0]["steps"] == 150
            for k in ("flat", "m", "v"):
                assert torch.equal(r[k], rs[0][k]), (mode, k)
            assert r["loss"] != rs[0]["loss"]          # each rank logs its own shard's loss
        assert bool(torch.isfinite(rs[0]["flat"]).all())
        res[mode] = rs[0]
    for k in ("flat", "m", "v"):
        if world == 2:      # a + b is one rounding in any order
            assert torch.equal(res["peer"][k], res["nccl"][k]), k
    # more ranks: NCCL's tree / ring order differs from the rank order, and 150 Adam steps (each ~lr in size whatever
    # the gradient's magnitude) amplify the reassociation noise: only the bulk has to agree
    assert float((res["peer"]["flat"] - res["nccl"]["flat"]).abs().mean()) < 1e-3
