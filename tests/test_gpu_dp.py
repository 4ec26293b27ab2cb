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
def test_two_rank_update_keeps_ranks_identical(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    out = str(tmp_path / "dp")
    res = {}
    for mode in ("peer", "nccl"):
        mp.spawn(_worker, args=(2, _free_port(), out, mode), nprocs=2, join=True)
        r0, r1 = torch.load(f"{out}.{mode}.0"), torch.load(f"{out}.{mode}.1")
        assert r0["steps"] == r1["steps"] == 150
        for k in ("flat", "m", "v"):
            assert torch.equal(r0[k], r1[k]), (mode, k)
        assert bool(torch.isfinite(r0["flat"]).all())
        assert r0["loss"] != r1["loss"]            # each rank logs its own shard's loss
        res[mode] = r0
    for k in ("flat", "m", "v"):
        assert torch.equal(res["peer"][k], res["nccl"][k]), k
