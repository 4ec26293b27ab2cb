"""GPU parity at the BASELINE.json configurations as WHOLE configs: one full rollout (act + scale + reset + synthetic
state + obs / reward / reset mask + id compaction, T steps), GAE over it, and one PPO minibatch step, each against the
CPU oracle driven the same way with the same explicit noise.

  configs[0]  18-DOF standing, N = 1000, T = 640, mini-chunk 40 (the reference formula, ppo.py:118-122)
  configs[2]  18-DOF walking + leg reward, 8192 envs per rank, T = 80, mini-chunk 5
  configs[3]  42-DOF joint set, N = 32768, T = 64, mini-chunk 4 — with A = 42 (O = 145) and with the reference's
              historical 18-of-42 actuation (fly.py:21-31, 272-292)
(configs[1], N = 16384 / T = 32, is the bench workload; its kernels are covered size by size in test_gpu_env.py /
test_gpu_agent.py and end to end by __graft_entry__.smoke().)

Bars: reset masks / ids bit-exact at every step; obs, reward, actions at 1e-5; GAE and the minibatch step as in
test_gpu_agent.py (1e-5 of scale for value-head quantities on the tensor-core path, 1e-4 for loss / gradients).
This is a CLOSED loop: the observation's commanded-action slots and the DOF targets are the policy's own actions
(held to 1e-5 relative / 2e-6 absolute in [-1, 1]) scaled to the joint range, whose half-width reaches 5.4 rad
(joint_LHTibia, assets/nmf_no_limits_limited_Dofs.urdf:227) — those entries therefore carry an absolute floor of
2e-6 x 5.4 ~ 1.1e-5, taken as 2e-5 here; every other entry keeps the open-loop bar of test_gpu_env.py.
The reward's joints-at-limit term is a STEP function of those commanded actions (cmd > 0.9 * upper, fly.py:736-737):
an action within its 2e-6 tolerance of a threshold may count one limit more or less, i.e. the reward moves by exactly
joints_at_limit_cost_scale = 0.1.  Such flips are allowed for at most 1e-5 of the rewards of a rollout (observed: 1-2
in 0.6-2 million) and must be whole multiples of 0.1; everything else is held to 1e-5.
"""
import pytest
import torch

import fly_bproject_b200 as fb
from oracle import fly_oracle as fo
from tests.helpers import ATOL_ENV, RTOL_ENV, RTOL_PPO, OraclePool, assert_obs_close, oracle_tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SCALED_ACTION_ATOL = 2e-5     # see the module docstring

CONFIGS = [
    pytest.param("limited_18", True, False, 1000, 640, 40, id="configs0-n1000-t640"),
    pytest.param("limited_18", True, True, 8192, 80, 5, id="configs2-walking-n8192-t80"),
    pytest.param("full_42", True, False, 32768, 64, 4, id="configs3-a42-n32768-t64"),
    pytest.param("full_42", False, False, 32768, 64, 4, id="configs3-a18of42-n32768-t64"),
]


@pytest.mark.timeout(900)
@pytest.mark.parametrize("asset,full,walking,n,horizon,mc", CONFIGS)
def test_whole_config_rollout_gae_and_one_minibatch_step(asset, full, walking, n, horizon, mc):
    t = oracle_tables(asset, n, full)
    a_dim, o_dim = t.num_act, t.num_obs
    torch.manual_seed(0)
    args = fb.make_args(num_envs=n, sim_device=DEV, horizon=horizon, verbose=False, state_pool=2, seed=4242,
                        asset=asset, full_actions=full, walking=walking)
    agent = fb.PPO(args)
    env = agent.env
    assert (agent.rollout_size, agent.mini_chunk_size, agent.num_obs, agent.num_acts) == (horizon, mc, o_dim, a_dim)
    if asset == "limited_18" and not walking:
        assert fo.rollout_shape(n) == (mc, horizon)          # the reference's own formula gives this config
    p0 = {k: v.cpu() for k, v in agent.net.state_dict().items()}

    s = fo.EnvState(t)
    pool = OraclePool(env.state)
    pool.bind(s)
    var = torch.full((a_dim,), 0.2)
    gen = torch.Generator().manual_seed(1)
    ring = torch.zeros(horizon + 1, n, o_dim)
    acts, logps = torch.zeros(horizon, n, a_dim), torch.zeros(horizon, n)
    rews = torch.zeros(horizon, n, 1)
    check_at = {0, 1, horizon // 2, horizon - 1}
    for i in range(horizon):
        noise = torch.randn(n, a_dim, generator=gen)
        ring[i] = s.obs_buf
        a, logp, _ = fo.act(p0, s.obs_buf, var, noise)                     # ppo.py:213-220
        acts[i], logps[i] = a, logp
        scaled, targets = fo.scale_actions(t, a)                           # fly.py:626-657
        s.actions = scaled
        fo.apply_reset(t, s)                                               # fly.py:446-480
        pool.simulate(s)
        fo.post_physics(t, s, walking=walking)                             # fly.py:676-681
        rews[i] = s.reward_buf.unsqueeze(-1)
        var = fo.anneal_var(var)
        agent._step_kernels(i, noise.to(DEV))
        assert torch.equal(env.reset_buf.cpu(), s.reset_buf), f"reset mask, step {i}"
        if i in check_at:
            want_ids = s.reset_buf.nonzero().squeeze(-1).to(torch.int32)
            cnt = int(env.reset_count.item())
            assert cnt == want_ids.numel() and torch.equal(env.reset_ids[:cnt].cpu(), want_ids)
            assert torch.equal(env.progress_buf.cpu(), s.progress_buf)
            torch.testing.assert_close(agent.all_acts[i].cpu(), a, rtol=1e-5, atol=2e-6)
            torch.testing.assert_close(agent.all_log_prob[i].cpu(), logp, rtol=1e-5, atol=1e-4)
            assert_obs_close(env.obs_buf.cpu(), s.obs_buf, a_dim, atol=SCALED_ACTION_ATOL)
            torch.testing.assert_close(env.dof_targets.cpu(), targets, rtol=1e-5, atol=SCALED_ACTION_ATOL)
            rerr = (env.reward_buf.cpu() - s.reward_buf).abs()
            roff = rerr > ATOL_ENV + RTOL_ENV * s.reward_buf.abs()
            assert int(roff.sum()) <= max(1, n // 100000), f"{int(roff.sum())} rewards off at step {i}"
    ring[horizon] = s.obs_buf
    torch.cuda.synchronize()
    torch.testing.assert_close(agent.action_var.cpu(), var, rtol=1e-6, atol=0)
    got_rew = agent.all_reward.cpu()
    err = (got_rew - rews).abs()
    off = err > ATOL_ENV + RTOL_ENV * rews.abs()
    assert float(off.float().mean()) <= 1e-5, f"{int(off.sum())} rewards off"
    if off.any():      # threshold flips of the joints-at-limit count: whole multiples of 0.1
        steps = err[off] / 0.1
        assert float((steps - steps.round()).abs().max()) < (0.2 if walking else 1e-3) and float(steps.round().min()) >= 1.0

    # ---- GAE over the whole rollout (ppo.py:157-171), on the ORACLE's buffers so that the comparison is of the scan
    #      and the value forward, not of accumulated rollout differences
    agent._obs_ring.copy_(ring)
    agent.all_acts.copy_(acts)
    agent.all_log_prob.copy_(logps)
    agent.all_reward.copy_(rews)
    done_last = (1 - s.reset_buf).unsqueeze(-1)
    target, adv = fo.make_data(p0, ring[:-1], ring[1:], rews, done_last)
    _, _, _, tg, ad = agent.make_data()
    vscale = float(target.abs().max())
    tc = agent.net.wpack is not None
    torch.testing.assert_close(tg.cpu(), target, rtol=1e-5, atol=max(2e-5, 1e-5 * vscale if tc else 0.0))
    torch.testing.assert_close(ad.cpu(), adv, rtol=1e-5, atol=max(5e-5, 2e-5 * float(adv.abs().max()) if tc else 0.0))

    # ---- one optimiser step on minibatch [0:mc] (ppo.py:183-199) from the oracle's target / advantage
    agent._target.copy_(target)
    agent.all_advantage.copy_(adv)
    p1 = {k: v.clone() for k, v in p0.items()}
    loss, norm, grads = fo.minibatch_step(p1, fo.AdamState(p1), ring[0:mc], acts[0:mc], logps[0:mc], target[0:mc],
                                          adv[0:mc], var)
    agent._minibatch(0, mc)
    torch.cuda.synchronize()
    assert float(agent.loss) == pytest.approx(loss, rel=RTOL_PPO, abs=1e-6)
    assert float(agent.grad_norm) == pytest.approx(norm, rel=RTOL_PPO)
    coef = min(1.0, 1.0 / (norm + 1e-6))
    gv, sd = agent.net.make_views(agent.grad), agent.net.state_dict()
    ill = 0
    for k in fo.PARAM_ORDER:
        scale = float(grads[k].abs().max()) * coef + 1e-12
        torch.testing.assert_close(gv[k].cpu(), grads[k] * coef, rtol=RTOL_PPO, atol=RTOL_PPO * scale, msg="grad " + k)
        well = grads[k].abs() > RTOL_PPO * float(grads[k].abs().max())     # Adam's first step: see test_gpu_agent.py
        err = (sd[k].cpu() - p1[k]).abs()
        tol = 1e-6 + RTOL_PPO * p1[k].abs()
        assert not ((err > tol) & well).any(), "param " + k
        assert float(err.max()) <= 2e-3 + 1e-6
        ill += int(((err > tol) & ~well).sum())
    assert ill <= 0.005 * agent.net.layout.total
