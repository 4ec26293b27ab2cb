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


# ---------------------------------------------------------------------------------------------------------------------
# The data-parallel CUDA step against the ORACLE: SURVEY.md §4 / §8(e) — "an N-shard step equals the reference
# single-process step on the concatenated batch up to fp reassociation" (reference ppo.py:183-199 is the step).
def _oracle_step_worker(rank, world, port, path, mode):
    import ctypes as C
    import torch.distributed as dist
    import fly_bproject_b200 as fb
    from fly_bproject_b200 import _lib, dp
    from oracle import fly_oracle as fo
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ["FLYB_DP"] = "peer"
    os.environ["FLYB_DP_TAIL"] = "0"
    dev = f"cuda:{rank}"
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    try:
        d = torch.load(path)
        o, a = d["obs"].shape[-1], d["acts"].shape[-1]
        net = fb.Net(o, a, dev, init=False, tensor_cores=True)
        net.load_state_dict(d["p0"])
        assert net.wpack is not None
        peers = dp.make_exchange(net.layout.total, torch.device(dev))
        assert peers is not None, "peer exchange unavailable on this box"
        n = d["obs"].shape[1]
        lo, hi = rank * n // world, (rank + 1) * n // world            # this rank's envs of the minibatch
        dv = [d[k][:, lo:hi].contiguous().to(dev) for k in ("obs", "acts", "logp", "target", "adv")]
        var = d["var"].to(dev)
        rows = dv[0].shape[0] * (hi - lo)
        lib, p = _lib.lib(), _lib.ptr
        cfg = _lib.PpoCfg(0.2, 1e-3, 0.9, 0.999, 1e-8, 1.0)
        ws = torch.zeros(lib.flyb_ppo_workspace_bytes(C.byref(net.layout), rows) + 256, dtype=torch.uint8, device=dev)
        wsp = (ws.data_ptr() + 255) // 256 * 256
        grad, m, v = torch.zeros_like(net.flat), torch.zeros_like(net.flat), torch.zeros_like(net.flat)
        loss, gn = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
        step = torch.zeros(1, dtype=torch.int64, device=dev)
        st = _lib.current_stream()
        if mode == "fused":
            _lib.check(lib.flyb_ppo_step_dp(C.byref(net.layout), C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]),
                                            p(dv[1]), p(dv[2]), p(dv[3]), p(dv[4]), p(var), rows, p(grad), p(loss), p(m),
                                            p(v), p(step), 1.0 / world, p(gn), wsp, peers.byref(), st), "flyb_ppo_step_dp")
        else:
            _lib.check(lib.flyb_ppo_grad(C.byref(net.layout), C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]), p(dv[1]),
                                         p(dv[2]), p(dv[3]), p(dv[4]), p(var), rows, p(grad), p(loss), wsp,
                                         peers.byref(), st), "flyb_ppo_grad")
            _lib.check(lib.flyb_ppo_apply(C.byref(net.layout), C.byref(cfg), p(net.flat), p(grad), p(m), p(v), p(step),
                                          1.0 / world, p(gn), wsp, p(net.wpack), peers.byref(), st), "flyb_ppo_apply")
        torch.cuda.synchronize()
        views = lambda flat: {k: x.cpu().clone() for k, x in net.make_views(flat).items()}
        torch.save({"w": {k: x.cpu() for k, x in net.state_dict().items()}, "m": views(m), "v": views(v),
                    "g": views(grad), "flat": net.flat.cpu(), "loss": float(loss), "norm": float(gn),
                    "step": int(step.item()), "status": peers.status()}, f"{path}.{mode}.{rank}")
        peers.close()
    finally:
        dist.destroy_process_group()


def _minibatch(shape):
    """(data dict, oracle results) of one optimiser step on the WHOLE minibatch."""
    from oracle import fly_oracle as fo
    from tests.helpers import load_golden
    if shape == "narrow":       # the golden minibatch of the executed reference: rows [0:mc] of ppo_n64_t32.npz
        g = load_golden("ppo_n64_t32.npz")
        mc = int(g["meta"][2])
        d = {"obs": g["roll_obs"][0:mc], "acts": g["roll_acts"][0:mc], "logp": g["roll_logp"][0:mc],
             "target": g["roll_target"][0:mc], "adv": g["roll_adv"][0:mc], "var": g["roll_var"],
             "p0": {k: g["w0/" + k].clone() for k in fo.PARAM_ORDER}}
    else:                       # the 42-joint shape (O = 145, A = 42): seeded synthetic minibatch, 2 x 512 rows
        o, a, t, n = 145, 42, 2, 512
        gen = torch.Generator().manual_seed(77)
        p0 = fo.init_params(o, a, seed=3)
        obs = torch.randn(t, n, o, generator=gen)
        var = torch.full((a,), 0.15)
        with torch.no_grad():
            mu = fo.pi(p0, obs.view(-1, o))
        act = (mu + 0.4 * torch.randn(t * n, a, generator=gen)).clip(-1, 1)
        old = fo.gaussian_logp(mu, act, var) + 0.3 * torch.randn(t * n, generator=gen)
        ratio = torch.exp(fo.gaussian_logp(mu, act, var) - old)        # keep rows clear of the clip's discontinuity
        old = torch.where(((ratio - 0.8).abs() < 1e-3) | ((ratio - 1.2).abs() < 1e-3), old + 0.05, old)
        d = {"obs": obs, "acts": act.view(t, n, a), "logp": old.view(t, n),
             "target": 2 * torch.randn(t, n, 1, generator=gen), "adv": torch.randn(t, n, 1, generator=gen), "var": var,
             "p0": p0}
    p1 = {k: x.clone() for k, x in d["p0"].items()}
    st = fo.AdamState(p1)
    o, a = d["obs"].shape[-1], d["acts"].shape[-1]
    loss, norm, grads = fo.minibatch_step(p1, st, d["obs"].reshape(-1, o), d["acts"].reshape(-1, a),
                                          d["logp"].reshape(-1), d["target"].reshape(-1, 1), d["adv"].reshape(-1, 1),
                                          d["var"])
    coef = min(1.0, 1.0 / (norm + 1e-6))
    return d, {"w": p1, "m": st.m, "v": st.v, "g": {k: x * coef for k, x in grads.items()}, "graw": grads,
               "loss": loss, "norm": norm}


@pytest.mark.timeout(300)
@pytest.mark.parametrize("shape,mode", [("narrow", "fused"), ("narrow", "split"), ("wide", "fused")])
def test_data_parallel_step_matches_oracle_on_the_whole_minibatch(tmp_path, shape, mode):
    """One flyb_ppo_step_dp (or flyb_ppo_grad + flyb_ppo_apply) per rank, each rank on ITS envs of a minibatch, against
    the oracle's optimiser step on the whole minibatch — for the narrow shape the golden step of the executed
    reference (w1 / m1 / v1 of ppo_n64_t32.npz).  Bars and the Adam-sign carve-out are those of the single-GPU test
    (test_gpu_agent.py: test_ppo_minibatch_step_matches_reference_golden): 1e-4."""
    from oracle import fly_oracle as fo
    from tests.helpers import RTOL_PPO, load_golden
    world = int(os.environ.get("FLYB_TEST_WORLD", "2"))
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    d, want = _minibatch(shape)
    if shape == "narrow":       # the oracle's step IS the reference's: pin it to the fixture before using it as the bar
        g = load_golden("ppo_n64_t32.npz")
        for k in fo.PARAM_ORDER:
            torch.testing.assert_close(want["w"][k], g["w1/" + k], rtol=1e-5, atol=1e-6)   # bar of test_oracle_golden.py
    path = str(tmp_path / f"mb_{shape}.pt")
    torch.save(d, path)
    _spawn(_oracle_step_worker, world, path, mode)
    rs = [torch.load(f"{path}.{mode}.{r}") for r in range(world)]
    for r in rs:
        assert r["status"] == 0 and r["step"] == 1
        assert torch.equal(r["flat"], rs[0]["flat"])                       # every rank holds the same parameters
        for k in fo.PARAM_ORDER:
            assert torch.equal(r["m"][k], rs[0]["m"][k]) and torch.equal(r["v"][k], rs[0]["v"][k]), k
    assert sum(r["loss"] for r in rs) / world == pytest.approx(want["loss"], rel=RTOL_PPO, abs=1e-6)
    got = rs[0]
    assert got["norm"] == pytest.approx(want["norm"], rel=1e-4)
    ill, total = 0, 0
    for k in fo.PARAM_ORDER:
        gref = want["g"][k]
        gmax = float(gref.abs().max())
        torch.testing.assert_close(got["g"][k], gref, rtol=RTOL_PPO, atol=RTOL_PPO * gmax, msg="clipped grad " + k)
        torch.testing.assert_close(got["m"][k], want["m"][k], rtol=RTOL_PPO,
                                   atol=RTOL_PPO * float(want["m"][k].abs().max()), msg="exp_avg " + k)
        torch.testing.assert_close(got["v"][k], want["v"][k], rtol=2 * RTOL_PPO,
                                   atol=2 * RTOL_PPO * float(want["v"][k].abs().max()), msg="exp_avg_sq " + k)
        # Adam's first step is +-lr wherever |g| >> eps: elements whose gradient is rounding noise of the row
        # reduction (below 1e-4 of the tensor's largest) may land anywhere within +-lr — counted and bounded
        well = gref.abs() > RTOL_PPO * gmax
        err = (got["w"][k] - want["w"][k]).abs()
        tol = 1e-6 + RTOL_PPO * want["w"][k].abs()
        bad = (err > tol) & well
        assert not bad.any(), f"param {k}: {int(bad.sum())} well-conditioned elements off, worst {float(err[well].max()):.3e}"
        assert float(err[~well].max() if (~well).any() else 0.0) <= 2.0 * 1e-3 + 1e-6, "param " + k
        ill += int(((err > tol) & ~well).sum())
        total += err.numel()
    assert ill <= 0.005 * total, f"{ill} near-zero-gradient elements moved differently"
