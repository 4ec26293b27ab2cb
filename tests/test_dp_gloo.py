"""Data-parallel step on CPU: world_size 2 over gloo.

The product kernels need a GPU; what runs here is the host-side sharding / collective plumbing (fly_bproject_b200/dp.py)
around the ORACLE's minibatch step: two ranks, each with its half of the env batch, all-reduce the gradient through
`dp.grad_all_reduce_` and must land on the parameters the reference's single-process step gives on the whole batch.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fly_bproject_b200 import dp
from oracle import fly_oracle as fo

O, A, MC, N = 73, 18, 2, 64


def _batch():
    g = torch.Generator().manual_seed(7)
    obs = torch.randn(MC, N, O, generator=g)
    act = torch.randn(MC, N, A, generator=g).clip(-1, 1)
    logp = -20 + torch.randn(MC, N, generator=g)
    target = torch.randn(MC, N, 1, generator=g)
    adv = torch.randn(MC, N, 1, generator=g)
    return obs, act, logp, target, adv


def _flat(d):
    return torch.cat([x.reshape(-1) for x in d.values()])


def _worker(rank, world_size, port, steps, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        torch.set_num_threads(1)
        assert dp.world() == (rank, world_size)
        start, cnt = dp.env_shard(N, rank, world_size)
        obs, act, logp, target, adv = (x[:, start:start + cnt] for x in _batch())
        p = fo.init_params(O, A, seed=rank)            # ranks start different: the broadcast must fix that
        flat = _flat(p)
        dp.broadcast_params_(flat, src=0)
        off = 0
        for k in p:
            n = p[k].numel()
            p[k] = flat[off:off + n].view_as(p[k]).clone()
            off += n
        st = fo.AdamState(p)
        var = torch.full((A,), 0.2)

        def hook(grads):
            flat_g = _flat(grads)
            scale = dp.grad_all_reduce_(flat_g)
            o2 = 0
            for k in grads:
                n = grads[k].numel()
                grads[k].copy_(flat_g[o2:o2 + n].view_as(grads[k]) * scale)
                o2 += n

        losses = []
        for _ in range(steps):
            loss, _, _ = fo.minibatch_step(p, st, obs, act, logp, target, adv, var, grad_hook=hook)
            losses.append(loss)
        # No CUDA device here: every rank fails to allocate its exchange buffer, the ranks agree on that collectively
        # (nobody is left waiting in a different mode) and the NCCL/gloo all-reduce path is selected.
        os.environ.pop("FLYB_DP", None)
        assert dp.exchange_mode() == "peer"
        assert dp.make_exchange(1000, torch.device("cpu")) is None
        os.environ["FLYB_DP"] = "nccl"
        assert dp.make_exchange(1000, torch.device("cpu")) is None
        torch.save({"p": p, "loss": losses}, f"{out_path}.{rank}")
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_env_shard_blocks():
    assert dp.env_shard(65536, 3, 8) == (3 * 8192, 8192)
    assert [dp.env_shard(16, r, 4) for r in range(4)] == [(0, 4), (4, 4), (8, 4), (12, 4)]
    with pytest.raises(ValueError):
        dp.env_shard(1000, 0, 3)
    with pytest.raises(ValueError):
        dp.env_shard(8, 2, 2)
    assert dp.world() == (0, 1)
    g = torch.ones(5)
    assert dp.grad_all_reduce_(g) == 1.0 and torch.equal(g, torch.ones(5))
    os.environ["FLYB_DP"] = "bogus"
    try:
        with pytest.raises(ValueError):
            dp.exchange_mode()
    finally:
        del os.environ["FLYB_DP"]


@pytest.mark.timeout(300)
def test_two_rank_step_equals_single_process_step(tmp_path):
    steps = 3
    out = str(tmp_path / "dp")
    mp.spawn(_worker, args=(2, _free_port(), steps, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    # every rank holds the same parameters after the steps
    for k in r0["p"]:
        assert torch.equal(r0["p"][k], r1["p"][k]), k
    # and they are the single-process step on the concatenated batch (reference ppo.py:183-199)
    p = fo.init_params(O, A, seed=0)
    st = fo.AdamState(p)
    var = torch.full((A,), 0.2)
    ref_losses = []
    for _ in range(steps):
        loss, _, _ = fo.minibatch_step(p, st, *_batch(), var)
        ref_losses.append(loss)
    for k in p:
        torch.testing.assert_close(r0["p"][k], p[k], rtol=1e-4, atol=1e-6, msg=k)
    # the logged loss of a rank is its shard's loss; the mean over ranks is the global loss
    for i in range(steps):
        assert abs(0.5 * (r0["loss"][i] + r1["loss"][i]) - ref_losses[i]) <= 1e-4 * max(1.0, abs(ref_losses[i]))
