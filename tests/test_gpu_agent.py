"""GPU parity, agent half: Net forward, act, GAE and the PPO minibatch step through the C ABI against the golden
vectors frozen from the executed reference (ppo.py) and against the CPU oracle.
Bar (north_star): GAE within 1e-5 relative; one PPO minibatch update within 1e-4 relative on loss and parameters."""
import ctypes as C
import math

import pytest
import torch

import fly_bproject_b200 as fb
from fly_bproject_b200 import _lib
from oracle import fly_oracle as fo
from tests.helpers import RTOL_PPO, OraclePool, load_golden, oracle_tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TC = True


@pytest.fixture(params=["tcgen05_3xtf32", "cuda_core_fp32"], autouse=True)
def gemm_path(request):
    """Every test runs on both GEMM paths: tensor cores (3xTF32, wpack given) and the exact-fp32 CUDA-core kernels."""
    global TC
    TC = request.param == "tcgen05_3xtf32"
    yield


def vclose(got, want, rtol, atol):
    """Value-head comparisons.  CUDA-core fp32 path: elementwise (rtol, atol).  Tensor-core path: a 3xTF32 product
    carries ~4e-7 * sum|a_k b_k| (hi/lo rounding; measured by tests/tc_probe), so outputs that are small next to
    their own intermediate sums — the reference checkpoint's value head reaches +-40 — are held to 1e-5 of the
    tensor's scale instead of 1e-5 of each element."""
    if TC:
        atol = max(atol, 1e-5 * float(want.abs().max()))
    torch.testing.assert_close(got, want, rtol=rtol, atol=atol)


def golden_net(g, tag="w0"):
    net = fb.Net(73, 18, DEV, init=False, tensor_cores=TC)
    net.load_state_dict({k: g[f"{tag}/{k}"] for k in fo.PARAM_ORDER})
    return net


def test_net_forward_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    net = golden_net(g)
    x = g["net_x"].to(DEV)
    torch.testing.assert_close(net.pi(x).cpu(), g["net_pi"], rtol=1e-5, atol=1e-5)
    vclose(net.v(x).cpu(), g["net_v"], rtol=1e-5, atol=2e-5)
    # checkpoint round trip in the reference's state_dict format (ppo.py:266-273)
    sd = net.state_dict()
    assert list(sd) == fo.PARAM_ORDER and sd["shared_net.0.weight"].shape == (256, 73)
    for k in fo.PARAM_ORDER:
        assert torch.equal(sd[k].cpu(), g["w0/" + k])


@pytest.mark.parametrize("o,a,rows", [(73, 18, 1), (73, 18, 16384 + 17), (145, 42, 3001), (145, 42, 1), (145, 42, 40000),
                                      (100, 27, 515)])
def test_net_forward_matches_oracle_random_weights(o, a, rows):
    p = fo.init_params(o, a, seed=o)
    net = fb.Net(o, a, DEV, init=False, tensor_cores=TC)
    assert (net.wpack is not None) == TC      # both shapes (O <= 80 / A <= 32 and O <= 160 / A <= 48) have a tcgen05 forward
    net.load_state_dict(p)
    x = torch.randn(rows, o, generator=torch.Generator().manual_seed(rows)) * 1.5
    mu, val = net._forward(x.to(DEV), True, True)
    torch.testing.assert_close(mu.cpu(), fo.pi(p, x), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(val.cpu(), fo.v(p, x), rtol=1e-5, atol=1e-5)


def _act(net, obs, var, noise, env=None, seed=0):
    rows = obs.shape[0]
    action = torch.empty(rows, net.num_act, device=DEV)
    logp = torch.empty(rows, device=DEV)
    mu = torch.empty(rows, net.num_act, device=DEV)
    p = _lib.ptr
    envargs = [None] * 9
    if env is not None:
        st = env.state
        envargs = [C.byref(env.cfg), p(env._actions_scaled), p(env.dof_targets), p(env.reset_buf), p(env.progress_buf),
                   p(env.potentials), p(env.prev_potentials), p(st.root), p(st.dof_state)]
    _lib.check(net.lib.flyb_policy_act(C.byref(net.layout), p(net.flat), p(net.wpack), p(obs), rows, p(var), p(noise), seed, 0, None,
                                       p(action), p(logp), p(mu), None, *envargs, _lib.current_stream()))
    return action, logp, mu


def test_act_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    net = golden_net(g)
    a, logp, _ = _act(net, g["act_obs"].to(DEV), g["act_var"].to(DEV), g["act_noise"].to(DEV))
    torch.testing.assert_close(a.cpu(), g["act_action"], rtol=1e-5, atol=2e-6)      # clipped action (ppo.py:220)
    torch.testing.assert_close(logp.cpu(), g["act_logp"], rtol=1e-5, atol=1e-5)     # of the UNclipped sample
    assert float(a.abs().max()) <= 1.0


def test_fused_act_pre_physics_equals_separate_kernels():
    n = 5000
    t = oracle_tables("full_42", n, False)
    st = fo.synth_state(t, seed=11)
    mk = lambda: fb.Fly(fb.make_args(num_envs=n, sim_device=DEV, asset="full_42", full_actions=False),
                        state_source=fb.TensorStateSource(st["root"].to(DEV), st["dof_states"].to(DEV),
                                                          st["force"].to(DEV)))
    e1, e2 = mk(), mk()
    mask = (torch.rand(n, generator=torch.Generator().manual_seed(1)) < 0.3).long().to(DEV)
    for e in (e1, e2):
        e.reset_buf.copy_(mask)
        e.progress_buf.fill_(7)
    net = fb.Net(73, 18, DEV, tensor_cores=TC)
    obs = torch.randn(n, 73, device=DEV)
    var = torch.full((18,), 0.2, device=DEV)
    noise = torch.randn(n, 18, device=DEV)
    a1, l1, _ = _act(net, obs, var, noise, env=e1)
    a2, l2, _ = _act(net, obs, var, noise)
    e2.pre_physics(a2)
    assert torch.equal(a1, a2) and torch.equal(l1, l2)
    for name in ["_actions_scaled", "dof_targets", "reset_buf", "progress_buf", "potentials", "prev_potentials"]:
        assert torch.equal(getattr(e1, name), getattr(e2, name)), name
    assert torch.equal(e1.state.root, e2.state.root) and torch.equal(e1.state.dof_state, e2.state.dof_state)
    # and against the oracle's scaling (fly.py:626-651)
    scaled, targets = fo.scale_actions(t, a1.cpu())
    torch.testing.assert_close(e1._actions_scaled.cpu(), scaled.view(n, -1), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(e1.dof_targets.cpu(), targets, rtol=1e-6, atol=1e-6)


def test_philox_noise_is_standard_normal_and_keyed():
    net = fb.Net(73, 18, DEV, tensor_cores=TC)
    n = 20000
    obs = torch.randn(n, 73, device=DEV)
    var = torch.full((18,), 0.04, device=DEV)
    a0, l0, mu0 = _act(net, obs, var, None, seed=5)
    a1, l1, _ = _act(net, obs, var, None, seed=5)
    a2, _, _ = _act(net, obs, var, None, seed=6)
    assert torch.equal(a0, a1) and torch.equal(l0, l1) and not torch.equal(a0, a2)
    inside = a0.abs() < 1.0
    eps = ((a0 - mu0) / 0.2)[inside]
    assert abs(float(eps.mean())) < 0.01 and abs(float(eps.std()) - 1.0) < 0.01
    # log-prob identity on rows where nothing was clipped
    rows = inside.all(dim=1)
    want = fo.gaussian_logp(mu0[rows].cpu(), a0[rows].cpu(), var.cpu())
    torch.testing.assert_close(l0[rows].cpu(), want, rtol=1e-4, atol=1e-3)


def _gae(reward, values, reset_last, gamma=0.99, lmbda=0.95):
    t, n = reward.shape[0], reward.shape[1]
    target = torch.empty(t, n, device=DEV)
    adv = torch.empty(t, n, device=DEV)
    p = _lib.ptr
    _lib.check(_lib.lib().flyb_gae(p(reward), p(values), p(reset_last), t, n, gamma, lmbda, p(target), p(adv),
                                   _lib.current_stream()))
    return target, adv


def test_gae_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    net = golden_net(g)
    ring = torch.cat([g["roll_obs"], g["roll_next"][-1:]], dim=0).to(DEV)      # (T+1, N, O)
    values = net.v(ring).squeeze(-1).contiguous()
    reset_last = (1 - g["roll_done"].squeeze(-1)).to(DEV)
    target, adv = _gae(g["roll_rew"].squeeze(-1).contiguous().to(DEV), values, reset_last)
    vclose(target.cpu(), g["roll_target"].squeeze(-1), rtol=1e-5, atol=2e-5)
    vclose(adv.cpu(), g["roll_adv"].squeeze(-1), rtol=1e-5, atol=5e-5)


@pytest.mark.parametrize("t,n", [(32, 16384), (1, 33), (7, 1000), (128, 4097)])
def test_gae_scan_bit_exact_vs_oracle_recursion_and_linear(t, n):
    gen = torch.Generator().manual_seed(t * n)
    reward = torch.randn(t, n, generator=gen)
    values = torch.randn(t + 1, n, generator=gen) * 3
    reset_last = (torch.rand(n, generator=gen) < 0.3).long()
    done = (1 - reset_last).float()
    # oracle recursion on the same values (ppo.py:160-169)
    target = reward + 0.99 * values[1:] * done
    delta = target - values[:-1]
    adv = torch.zeros_like(delta)
    run = torch.zeros(n)
    gl = torch.tensor(0.99 * 0.95, dtype=torch.float32)
    for i in range(t - 1, -1, -1):
        run = gl * run + delta[i]
        adv[i] = run
    tg, ad = _gae(reward.to(DEV), values.to(DEV), reset_last.to(DEV))
    assert torch.equal(tg.cpu(), target)
    torch.testing.assert_close(ad.cpu(), adv, rtol=1e-6, atol=1e-6)
    # size-independent property: the scan is linear in (reward, values)
    tg2, ad2 = _gae((2 * reward).to(DEV), (2 * values).to(DEV), reset_last.to(DEV))
    torch.testing.assert_close(ad2, 2 * ad, rtol=1e-5, atol=1e-5)


def _grad_step(net, g, k, j, var):
    """One flyb_ppo_grad on golden rollout rows [k:j]; returns (grad views, loss)."""
    n = g["roll_obs"].shape[1]
    rows = (j - k) * n
    lib, p = _lib.lib(), _lib.ptr
    cfg = _lib.PpoCfg(0.2, 1e-3, 0.9, 0.999, 1e-8, 1.0)
    ws = torch.zeros(lib.flyb_ppo_workspace_bytes(C.byref(net.layout), rows) + 256, dtype=torch.uint8, device=DEV)
    wsp = (ws.data_ptr() + 255) // 256 * 256
    grad = torch.zeros_like(net.flat)
    loss = torch.zeros(1, device=DEV)
    # keep the device copies alive until the launch has finished
    dv = [x.contiguous().to(DEV) for x in (g["roll_obs"][k:j], g["roll_acts"][k:j], g["roll_logp"][k:j],
                                           g["roll_target"][k:j], g["roll_adv"][k:j], var)]
    _lib.check(lib.flyb_ppo_grad(C.byref(net.layout), C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]), p(dv[1]), p(dv[2]),
                                 p(dv[3]), p(dv[4]), p(dv[5]), rows, p(grad), p(loss), wsp, None, _lib.current_stream()))
    torch.cuda.synchronize()
    return grad, loss, ws, wsp, cfg


def test_ppo_minibatch_step_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    net = golden_net(g)
    grad, loss, ws, wsp, cfg = _grad_step(net, g, 0, mc, g["roll_var"])
    torch.testing.assert_close(loss.cpu().squeeze(), g["loss0"], rtol=RTOL_PPO, atol=1e-6)
    # unclipped gradient against the oracle's autograd on the same rows
    p0 = {k: g["w0/" + k].clone() for k in fo.PARAM_ORDER}
    _, norm, grads = fo.minibatch_step(p0, fo.AdamState(p0), g["roll_obs"][0:mc], g["roll_acts"][0:mc],
                                       g["roll_logp"][0:mc], g["roll_target"][0:mc], g["roll_adv"][0:mc], g["roll_var"])
    gv = net.make_views(grad)
    for k in fo.PARAM_ORDER:
        scale = float(grads[k].abs().max()) + 1e-12
        torch.testing.assert_close(gv[k].cpu(), grads[k], rtol=RTOL_PPO, atol=RTOL_PPO * scale, msg="grad " + k)
    # padding of the flat layout carries exactly zero gradient
    assert float(grad.abs().sum()) == pytest.approx(float(sum(gv[k].abs().sum() for k in fo.PARAM_ORDER)), rel=1e-6)
    # clip + Adam against the golden parameters / moments after optimiser step 1 (ppo.py:198-199)
    lib, p = _lib.lib(), _lib.ptr
    m, v = torch.zeros_like(net.flat), torch.zeros_like(net.flat)
    step = torch.zeros(1, dtype=torch.int64, device=DEV)
    gn = torch.zeros(1, device=DEV)
    _lib.check(lib.flyb_ppo_apply(C.byref(net.layout), C.byref(cfg), p(net.flat), p(grad), p(m), p(v), p(step), 1.0,
                                  p(gn), wsp, p(net.wpack), None, _lib.current_stream()))
    torch.cuda.synchronize()
    assert int(step.item()) == 1
    assert float(gn) == pytest.approx(norm, rel=1e-4)
    sd, mv, vv, gc = net.state_dict(), net.make_views(m), net.make_views(v), net.make_views(grad)
    ill = 0
    for k in fo.PARAM_ORDER:
        # Adam's first step is lr * g / (|g| + eps) = +-lr wherever |g| >> eps: it is discontinuous in g at 0, so an
        # element whose gradient is below the gradient tolerance band (1e-4 of the tensor's largest gradient — i.e.
        # rounding noise of the reduction over rows) may legitimately land anywhere within +-lr of the reference.
        gref = g["g0/" + k]
        band = RTOL_PPO * float(gref.abs().max())
        well = gref.abs() > band
        got, want = sd[k].cpu(), g["w1/" + k]
        err = (got - want).abs()
        tol = 1e-6 + RTOL_PPO * want.abs()
        bad = (err > tol) & well
        assert not bad.any(), f"param {k}: {int(bad.sum())} well-conditioned elements off, worst {float(err[well].max()):.3e}"
        assert float(err[~well].max() if (~well).any() else 0.0) <= 2.0 * 1e-3 + 1e-6, "param " + k
        ill += int(((err > tol) & ~well).sum())
        torch.testing.assert_close(gc[k].cpu(), g["g0/" + k], rtol=RTOL_PPO,
                                   atol=RTOL_PPO * float(g["g0/" + k].abs().max()), msg="clipped grad " + k)
        torch.testing.assert_close(mv[k].cpu(), g["m1/" + k], rtol=RTOL_PPO,
                                   atol=RTOL_PPO * float(g["m1/" + k].abs().max()), msg="exp_avg " + k)
        torch.testing.assert_close(vv[k].cpu(), g["v1/" + k], rtol=2 * RTOL_PPO,
                                   atol=2 * RTOL_PPO * float(g["v1/" + k].abs().max()), msg="exp_avg_sq " + k)
    assert ill <= 0.005 * 69587, f"{ill} near-zero-gradient elements moved differently"


class _ReplayEnv:
    """Foreign env behind the reference's duck-typed surface: replays the golden rollout."""

    def __init__(self, g):
        self.g, self.i = g, 0
        self.num_act, self.num_obs, self.end = 18, 73, False
        self.obs_buf = g["roll_obs"][0].to(DEV)
        self.reward_buf = torch.zeros(64, device=DEV)
        self.reset_buf = torch.ones(64, dtype=torch.long, device=DEV)

    def step(self, a):
        g, i = self.g, self.i
        self.obs_buf = g["roll_next"][i].to(DEV)
        self.reward_buf = g["roll_rew"][i].squeeze(-1).to(DEV)
        self.reset_buf = (1 - g["roll_done"].squeeze(-1)).to(DEV) if i == g["roll_obs"].shape[0] - 1 else \
            torch.zeros(64, dtype=torch.long, device=DEV)
        self.i += 1


def test_full_update_matches_reference_golden_75_steps():
    """PPO.update through the drop-in class on the golden rollout: 5 epochs x 15 minibatches (log.txt:49)."""
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    args = fb.make_args(num_envs=n, sim_device=DEV, horizon=t_roll, verbose=False, tensor_cores=TC)
    agent = fb.PPO(args, env=_ReplayEnv(g))
    assert agent.mini_chunk_size == mc and agent.rollout_size == t_roll
    agent.net.load_state_dict({k: g["w0/" + k] for k in fo.PARAM_ORDER})
    agent._obs_ring[:t_roll].copy_(g["roll_obs"])
    agent._obs_ring[t_roll].copy_(g["roll_next"][-1])
    agent.all_acts.copy_(g["roll_acts"])
    agent.all_reward.copy_(g["roll_rew"])
    agent.all_log_prob.copy_(g["roll_logp"])
    agent.action_var.copy_(g["roll_var"])
    agent.env.reset_buf = (1 - g["roll_done"].squeeze(-1)).to(DEV)
    assert agent.all_done.shape == (n, 1) and agent.all_done.dtype == torch.int64       # ppo.py:230 quirk
    obs, acts, logp, target, adv = agent.make_data()
    vclose(target.cpu(), g["roll_target"], rtol=1e-5, atol=2e-5)
    vclose(adv.cpu(), g["roll_adv"], rtol=1e-5, atol=5e-5)
    agent.update()
    assert agent.optim_step == 75 and int(agent.adam_step.item()) == 75
    sd = agent.net.state_dict()
    for k in fo.PARAM_ORDER:
        torch.testing.assert_close(sd[k].cpu(), g["w75/" + k], rtol=2e-3, atol=2e-4, msg="param " + k)


def test_three_minibatch_steps_match_reference_golden():
    """Optimiser steps 1-3 of the reference update (minibatches [0:mc], [mc:2mc], [2mc:3mc] of the golden rollout,
    ppo.py:178-202) through flyb_ppo_step, against the frozen w3 / m3 / v3 at 1e-4.  Same carve-out as the one-step
    test: Adam moves a parameter by ~lr * sign(g) while |g| >> eps, so elements whose gradient sat inside the rounding
    band (1e-4 of the tensor's largest gradient) in ANY of the three steps may land up to 3 * lr away; they are counted
    and bounded, every other element is held to 1e-4."""
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    net = golden_net(g)
    lib, p = _lib.lib(), _lib.ptr
    cfg = _lib.PpoCfg(0.2, 1e-3, 0.9, 0.999, 1e-8, 1.0)
    rows = mc * n
    ws = torch.zeros(lib.flyb_ppo_workspace_bytes(C.byref(net.layout), rows) + 256, dtype=torch.uint8, device=DEV)
    wsp = (ws.data_ptr() + 255) // 256 * 256
    grad, loss = torch.zeros_like(net.flat), torch.zeros(1, device=DEV)
    m, v = torch.zeros_like(net.flat), torch.zeros_like(net.flat)
    step, gn = torch.zeros(1, dtype=torch.int64, device=DEV), torch.zeros(1, device=DEV)
    var = g["roll_var"].to(DEV)
    # oracle on the same three minibatches: the well-conditioned mask of every step, and a cross-check of w3 itself
    p_or = {k: g["w0/" + k].clone() for k in fo.PARAM_ORDER}
    st_or = fo.AdamState(p_or)
    well = {k: torch.ones_like(p_or[k], dtype=torch.bool) for k in fo.PARAM_ORDER}
    keep = []
    for i in range(3):
        k0, k1 = i * mc, (i + 1) * mc
        dv = [x.contiguous().to(DEV) for x in (g["roll_obs"][k0:k1], g["roll_acts"][k0:k1], g["roll_logp"][k0:k1],
                                               g["roll_target"][k0:k1], g["roll_adv"][k0:k1])]
        keep.append(dv)
        _lib.check(lib.flyb_ppo_step(C.byref(net.layout), C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]), p(dv[1]),
                                     p(dv[2]), p(dv[3]), p(dv[4]), p(var), rows, p(grad), p(loss), p(m), p(v), p(step),
                                     p(gn), wsp, _lib.current_stream()))
        _, _, grads = fo.minibatch_step(p_or, st_or, g["roll_obs"][k0:k1], g["roll_acts"][k0:k1], g["roll_logp"][k0:k1],
                                        g["roll_target"][k0:k1], g["roll_adv"][k0:k1], g["roll_var"])
        for k in fo.PARAM_ORDER:
            well[k] &= grads[k].abs() > RTOL_PPO * float(grads[k].abs().max())
    torch.cuda.synchronize()
    assert int(step.item()) == 3
    sd, mv, vv = net.state_dict(), net.make_views(m), net.make_views(v)
    ill = 0
    for k in fo.PARAM_ORDER:
        torch.testing.assert_close(p_or[k], g["w3/" + k], rtol=1e-5, atol=1e-6, msg="oracle w3 " + k)
        got, want = sd[k].cpu(), g["w3/" + k]
        err = (got - want).abs()
        tol = 1e-6 + RTOL_PPO * want.abs()
        bad = (err > tol) & well[k]
        assert not bad.any(), f"param {k}: {int(bad.sum())} well-conditioned elements off, worst {float(err[well[k]].max()):.3e}"
        assert float(err.max()) <= 3.0 * 2.0 * 1e-3 + 1e-6, "param " + k
        ill += int(((err > tol) & ~well[k]).sum())
        torch.testing.assert_close(mv[k].cpu(), g["m3/" + k], rtol=RTOL_PPO,
                                   atol=RTOL_PPO * float(g["m3/" + k].abs().max()), msg="exp_avg " + k)
        torch.testing.assert_close(vv[k].cpu(), g["v3/" + k], rtol=2 * RTOL_PPO,
                                   atol=2 * RTOL_PPO * float(g["v3/" + k].abs().max()), msg="exp_avg_sq " + k)
    assert ill <= 0.01 * 69587, f"{ill} near-zero-gradient elements moved differently"


@pytest.mark.parametrize("o,a,rows", [(145, 42, 1000), (145, 42, 32768 + 5), (145, 42, 7), (73, 18, 32768), (73, 18, 100),
                                      (100, 27, 300)])
def test_ppo_grad_matches_oracle_autograd_other_shapes(o, a, rows):
    gen = torch.Generator().manual_seed(rows + o)
    p0 = fo.init_params(o, a, seed=3)
    net = fb.Net(o, a, DEV, init=False, tensor_cores=TC)
    net.load_state_dict(p0)
    obs = torch.randn(rows, o, generator=gen)
    var = torch.full((a,), 0.15)
    with torch.no_grad():
        mu = fo.pi(p0, obs)
    act = (mu + 0.4 * torch.randn(rows, a, generator=gen)).clip(-1, 1)
    old_logp = fo.gaussian_logp(mu, act, var) + 0.3 * torch.randn(rows, generator=gen)   # ratios in and out of the clip
    # the clipped surrogate's gradient is DISCONTINUOUS where the ratio crosses 1 +- clip: a row within rounding distance
    # of a boundary contributes its whole gradient or nothing (a 1/sqrt(rows) share of a cancelling sum).  Keep every row
    # a clear 1e-3 away so that the comparison is of the arithmetic, not of a coin flip.
    ratio = torch.exp(fo.gaussian_logp(mu, act, var) - old_logp)
    near = ((ratio - 0.8).abs() < 1e-3) | ((ratio - 1.2).abs() < 1e-3)
    old_logp = torch.where(near, old_logp + 0.05, old_logp)
    target = torch.randn(rows, 1, generator=gen) * 2
    adv = torch.randn(rows, 1, generator=gen)
    g = {"roll_obs": obs.view(1, rows, o), "roll_acts": act.view(1, rows, a), "roll_logp": old_logp.view(1, rows),
         "roll_target": target.view(1, rows, 1), "roll_adv": adv.view(1, rows, 1)}
    grad, loss, *_ = _grad_step(net, g, 0, 1, var)
    leaf = {k: x.clone().requires_grad_(True) for k, x in p0.items()}
    want = fo.minibatch_loss(leaf, obs, act, old_logp, target, adv, var)
    want.backward()
    torch.testing.assert_close(loss.cpu().squeeze(), want.detach(), rtol=RTOL_PPO, atol=1e-6)
    gv = net.make_views(grad)
    for k in fo.PARAM_ORDER:
        scale = float(leaf[k].grad.abs().max()) + 1e-12
        torch.testing.assert_close(gv[k].cpu(), leaf[k].grad, rtol=RTOL_PPO, atol=RTOL_PPO * scale, msg="grad " + k)


def test_run_loop_matches_oracle_rollout_and_update():
    """Integration: the drop-in PPO + Fly driven like trainer.py, against the oracle driven the same way, with
    explicit noise; two env steps then compare rollout rows."""
    n = 512
    t = oracle_tables("limited_18", n, True)
    args = fb.make_args(num_envs=n, sim_device=DEV, horizon=16, verbose=False, state_pool=2, seed=77, tensor_cores=TC)
    agent = fb.PPO(args)
    env = agent.env
    p0 = {k: v.cpu() for k, v in agent.net.state_dict().items()}
    # oracle env on the same synthetic pool
    s = fo.EnvState(t)
    pool = OraclePool(env.state)
    var = torch.full((18,), 0.2)
    pool.bind(s)
    gen = torch.Generator().manual_seed(0)
    for i in range(3):
        noise = torch.randn(n, 18, generator=gen)
        obs_before = s.obs_buf.clone()
        a, logp, _ = fo.act(p0, obs_before, var, noise)
        scaled, targets = fo.scale_actions(t, a)
        s.actions = scaled
        fo.apply_reset(t, s)
        pool.simulate(s)
        fo.post_physics(t, s)
        var = fo.anneal_var(var)
        agent.run(noise=noise.to(DEV))
        torch.cuda.synchronize()
        torch.testing.assert_close(agent.all_acts[i].cpu(), a, rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(agent.all_log_prob[i].cpu(), logp, rtol=1e-5, atol=1e-4)
        torch.testing.assert_close(agent.all_obs[i].cpu(), obs_before, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(agent.all_reward[i].squeeze(-1).cpu(), s.reward_buf, rtol=1e-5, atol=1e-5)
        assert torch.equal(env.reset_buf.cpu(), s.reset_buf)
        torch.testing.assert_close(agent.action_var.cpu(), var, rtol=1e-6, atol=0)
    assert agent.mini_batch_number == 3 and agent.run_step == 3


def test_ppo_grad_is_deterministic():
    """Fixed-order partial reduction, no float atomics: the same minibatch gives bit-identical gradients and loss."""
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    net = golden_net(g)
    big = {k: (v.repeat(1, 40, 1) if v.dim() == 3 else v.repeat(1, 40)) for k, v in g.items()
           if k in ("roll_obs", "roll_acts", "roll_logp", "roll_target", "roll_adv")}
    outs = []
    for _ in range(3):
        grad, loss, *_ = _grad_step(net, big, 0, mc, g["roll_var"])
        outs.append((grad.clone(), loss.clone()))
    for grad, loss in outs[1:]:
        assert torch.equal(grad, outs[0][0]) and torch.equal(loss, outs[0][1])


def _fresh_pack(net):
    """A second weight stream packed from scratch from the net's CURRENT parameters."""
    lib = _lib.lib()
    buf = torch.zeros_like(net.wpack)
    _lib.check(lib.flyb_pack_weights(C.byref(net.layout), _lib.ptr(net.flat), _lib.ptr(buf), _lib.current_stream()))
    torch.cuda.synchronize()
    return buf


def test_ppo_step_is_grad_plus_apply_and_keeps_the_weight_stream_in_step():
    """flyb_ppo_step (norm taken by the gradient reduction, weight stream refreshed in place by the Adam kernel)
    against flyb_ppo_grad + flyb_ppo_apply on the same minibatch, three optimiser steps in a row; after every step the
    in-place refreshed stream must equal a from-scratch flyb_pack_weights of the new parameters bit for bit."""
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    lib, p = _lib.lib(), _lib.ptr
    cfg = _lib.PpoCfg(0.2, 1e-3, 0.9, 0.999, 1e-8, 1.0)
    rows = mc * n
    dv = [x.contiguous().to(DEV) for x in (g["roll_obs"][0:mc], g["roll_acts"][0:mc], g["roll_logp"][0:mc],
                                           g["roll_target"][0:mc], g["roll_adv"][0:mc], g["roll_var"])]
    out = {}
    import os
    for mode in ("split", "fused", "fused_3", "fused_tail"):
        # "fused": four launches (the default); "fused_3": FLYB_ADAM_FUSED=1, clip + Adam inside the gradient reduction
        # (three launches; opt-in, measured slower); "fused_tail": FLYB_PPO_TAIL=1, the two-launch form (reduction + clip + Adam behind grid
        # barriers in the tail of the cooperatively launched weight-gradient kernel) — opt-in, measured slower
        os.environ["FLYB_PPO_TAIL"] = "1" if mode == "fused_tail" else "0"
        os.environ["FLYB_ADAM_FUSED"] = "1" if mode == "fused_3" else "0"
        net = golden_net(g)
        ws = torch.zeros(lib.flyb_ppo_workspace_bytes(C.byref(net.layout), rows) + 256, dtype=torch.uint8, device=DEV)
        wsp = (ws.data_ptr() + 255) // 256 * 256
        grad, loss = torch.zeros_like(net.flat), torch.zeros(1, device=DEV)
        m, v = torch.zeros_like(net.flat), torch.zeros_like(net.flat)
        step, gn = torch.zeros(1, dtype=torch.int64, device=DEV), torch.zeros(1, device=DEV)
        lay, st = C.byref(net.layout), _lib.current_stream()
        norms = []
        for it in range(3):
            if mode == "split":
                _lib.check(lib.flyb_ppo_grad(lay, C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]), p(dv[1]), p(dv[2]),
                                             p(dv[3]), p(dv[4]), p(dv[5]), rows, p(grad), p(loss), wsp, None, st))
                _lib.check(lib.flyb_ppo_apply(lay, C.byref(cfg), p(net.flat), p(grad), p(m), p(v), p(step), 1.0, p(gn),
                                              wsp, p(net.wpack), None, st))
            else:
                _lib.check(lib.flyb_ppo_step(lay, C.byref(cfg), p(net.flat), p(net.wpack), p(dv[0]), p(dv[1]), p(dv[2]),
                                             p(dv[3]), p(dv[4]), p(dv[5]), rows, p(grad), p(loss), p(m), p(v), p(step),
                                             p(gn), wsp, st))
            torch.cuda.synchronize()
            assert int(step.item()) == it + 1
            if net.wpack is not None:   # (the CUDA-core variant of this test has no weight stream)
                slabs = net.wpack.numel() - net.layout.total * 16      # behind them: the parameter -> slab index
                assert torch.equal(net.wpack[:slabs], _fresh_pack(net)[:slabs]), \
                    f"{mode}: weight stream out of step after step {it + 1}"
            norms.append(float(gn))
            if it == 0:
                out[mode] = [net.flat.clone(), m.clone(), v.clone(), loss.clone()]
        out[mode] += [float(loss), norms, net.flat.clone(), m.clone(), v.clone()]
    os.environ.pop("FLYB_PPO_TAIL", None)
    os.environ.pop("FLYB_ADAM_FUSED", None)
    if TC:   # Adam inside the reduction adds the norm partials in the Adam kernel's own order: nothing moves, not a bit
        a, b = out["fused"], out["fused_3"]
        assert a[4] == b[4] and a[5] == b[5]
        for k in (0, 1, 2, 3, 6, 7, 8):
            assert torch.equal(a[k], b[k]), k
    a = out["split"]
    for other in ("fused", "fused_3", "fused_tail"):
        b = out[other]
        # step 1: same gradient bit for bit; the forms differ only in how the norm's partial sums are grouped, i.e. the
        # clip coefficient moves by an ulp or two.  Later steps start from those ulp-apart parameters.
        assert torch.equal(a[3], b[3]), other
        for x, y, name in zip(a[:3], b[:3], ("params", "exp_avg", "exp_avg_sq")):
            torch.testing.assert_close(x, y, rtol=1e-5, atol=1e-7, msg=f"{other} {name}")
        assert a[4] == pytest.approx(b[4], rel=1e-4)
        assert a[5] == pytest.approx(b[5], rel=1e-4)


@pytest.mark.parametrize("n", [1001, 301])
def test_env_counts_that_misalign_the_rollout_rows(n):
    """N * O not a multiple of 4 (N = 1001 or 301, O = 73): the rollout-ring rows obs_copy / all_acts[i] are then only
    4- or 8-byte aligned.  The reference accepts any N; the kernels must fall back to narrower stores."""
    t = oracle_tables("limited_18", n, True)
    args = fb.make_args(num_envs=n, sim_device=DEV, horizon=16, verbose=False, state_pool=2, seed=n, tensor_cores=TC)
    agent = fb.PPO(args)
    env = agent.env
    p0 = {k: v.cpu() for k, v in agent.net.state_dict().items()}
    s = fo.EnvState(t)
    pool = OraclePool(env.state)
    pool.bind(s)
    var = torch.full((18,), 0.2)
    gen = torch.Generator().manual_seed(n)
    for i in range(5):
        noise = torch.randn(n, 18, generator=gen)
        a, logp, _ = fo.act(p0, s.obs_buf.clone(), var, noise)
        s.actions, _ = fo.scale_actions(t, a)
        fo.apply_reset(t, s)
        pool.simulate(s)
        fo.post_physics(t, s)
        var = fo.anneal_var(var)
        agent.run(noise=noise.to(DEV))
        torch.cuda.synchronize()
        assert torch.equal(env.reset_buf.cpu(), s.reset_buf)
        assert torch.equal(agent._obs_ring[i + 1], env.obs_buf)            # the (possibly misaligned) ring copy
        torch.testing.assert_close(agent.all_acts[i].cpu(), a, rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(env._actions_scaled.cpu(), s.actions.view(n, -1), rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(agent.all_reward[i].squeeze(-1).cpu(), s.reward_buf, rtol=1e-5, atol=1e-5)
    for i in range(5, 16):
        agent.run()                                                        # the 16th step triggers the update
    torch.cuda.synchronize()
    assert agent.optim_step == 75 and bool(torch.isfinite(agent.net.flat).all())


@pytest.mark.parametrize("n", [16384, 100, 4096 + 77, 128 * 148, 128 * 148 - 5, 40000, 1])
def test_fused_step_compacts_its_own_reset_ids(n):
    """The id list the fused env step leaves behind (every tile writes its own ids once the tiles in front have published
    their counts) == the envs flagged by that step (reset_buf.nonzero() at fly.py:447), ascending, with the count — checked on the LAST fused launch of a
    rollout, i.e. with the closing obs / reward / reset launch (which would rewrite the list) left out; repeated, so that
    the launch ticket and the counter re-arm are exercised across launches and across the two kernels."""
    if not TC:
        pytest.skip("the fused step exists on the tensor-core path only")
    horizon = 6
    ag = fb.PPO(fb.make_args(num_envs=n, sim_device=DEV, horizon=horizon, verbose=False, state_pool=3, seed=5,
                             fused_steps=True))
    env = ag.env
    closing = env.post_physics
    seen = set()
    for rep in range(3):
        env.post_physics = lambda **kw: None          # rollout without its closing launch
        try:
            ag._rollout_kernels()
        finally:
            env.post_physics = closing
        torch.cuda.synchronize()
        # the launch's act half has already performed the reset and cleared the mask (fly.py:476-477): the envs it listed
        # are the ones whose progress counter is back at 0 (everyone else's was incremented, fly.py:678)
        want = (env.progress_buf == 0).nonzero().squeeze(-1).to(torch.int32)
        assert int(env.reset_buf.sum()) == 0
        cnt = int(env.reset_count.item())
        assert cnt == want.numel(), (rep, cnt, want.numel())
        assert torch.equal(env.reset_ids[:cnt], want), rep
        seen.add(cnt)
        if rep == 1:                                   # a stand-alone launch in between shares the counter
            env.post_physics()
            torch.cuda.synchronize()
            want = env.reset_buf.nonzero().squeeze(-1).to(torch.int32)
            assert int(env.reset_count.item()) == want.numel()
            assert torch.equal(env.reset_ids[:want.numel()], want)
    assert n < 64 or max(seen) > 0


@pytest.mark.parametrize("n,horizon,pool,explicit_noise", [(16384, 32, 16, False), (40000, 8, 4, True), (100, 4, 2, False),
                                                           (4096 + 77, 16, 4, True)])
def test_fused_rollout_is_bit_identical_to_the_two_kernel_rollout(n, horizon, pool, explicit_noise):
    """flyb_env_step_fused (one launch per env step: obs + reward + reset of step i-1, act + pre-physics of step i) against
    the two-launch schedule on the same seeds: every rollout buffer, env buffer, the synthetic state (reset rows), the
    compacted ids, the annealed variance and the Philox counter must agree bit for bit.  Sizes cover one tile per CTA,
    the persistent multi-tile loop with a ragged last tile, and a single ragged tile."""
    if not TC:
        pytest.skip("the fused step exists on the tensor-core path only")
    agents = []
    noise = None
    if explicit_noise:
        noise = torch.randn(horizon, n, 18, generator=torch.Generator().manual_seed(n)).to(DEV)
    for fused in (True, False):
        torch.manual_seed(3)
        args = fb.make_args(num_envs=n, sim_device=DEV, horizon=horizon, verbose=False, state_pool=pool, seed=11,
                            fused_steps=fused)
        ag = fb.PPO(args)
        assert ag._fused_steps == fused
        for _ in range(2):                       # the second rollout starts from the first one's closing state
            ag.collect_rollout(use_graph=False, noise=noise)
        agents.append(ag)
    torch.cuda.synchronize()
    a, b = agents
    for name in ("_obs_ring", "all_acts", "all_log_prob", "all_reward", "action_var", "_rng_counter"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    for name in ("obs_buf", "reward_buf", "reset_buf", "progress_buf", "potentials", "prev_potentials", "up_vec",
                 "heading_vec", "_actions_scaled", "dof_targets", "reset_count"):
        assert torch.equal(getattr(a.env, name), getattr(b.env, name)), name
    cnt = int(a.env.reset_count.item())
    assert torch.equal(a.env.reset_ids[:cnt], b.env.reset_ids[:cnt])
    assert torch.equal(a.env.reset_ids[:cnt], a.env.reset_buf.nonzero().squeeze(-1).to(torch.int32))
    for ea, eb in zip(a.env.state.pool, b.env.state.pool):
        assert torch.equal(ea["root"], eb["root"]) and torch.equal(ea["dof_state"], eb["dof_state"])
    torch.testing.assert_close(a.env.reward_sum, b.env.reward_sum, rtol=1e-5, atol=1e-2)
    assert int(a._rng_counter.item()) == 2 * horizon
    # and through a captured graph (what bench.py replays): same as eager, fused against fused
    torch.manual_seed(3)
    c = fb.PPO(fb.make_args(num_envs=n, sim_device=DEV, horizon=horizon, verbose=False, state_pool=pool, seed=11))
    if horizon % pool == 0 and not explicit_noise:
        c.collect_rollout(use_graph=True)        # warm-up rollout + capture
        c.collect_rollout(use_graph=True)        # replay
        torch.cuda.synchronize()
        for name in ("_obs_ring", "all_acts", "all_log_prob", "all_reward", "action_var"):
            assert torch.equal(getattr(a, name), getattr(c, name)), name


def test_captured_rollout_refuses_explicit_noise():
    """A captured rollout draws from Philox: handing it a noise tensor must fail loudly, not be dropped."""
    ag = fb.PPO(fb.make_args(num_envs=256, sim_device=DEV, horizon=4, verbose=False, state_pool=2))
    noise = torch.zeros(4, 256, 18, device=DEV)
    with pytest.raises(fb.FlybError, match="use_graph=False"):
        ag.collect_rollout(noise=noise)
    ag.collect_rollout(use_graph=False, noise=noise)        # eager takes it: zero noise -> action == clip(mu)
    assert ag.run_step == 4
