"""GPU parity, env half: the CUDA kernels called through the C ABI (via the `Fly` drop-in) against
(1) golden vectors frozen from the executed reference and (2) the CPU oracle on seeded synthetic states.
Bar (north_star): reset masks / ids / counters bit-exact; obs and reward within 1e-5 relative."""
import pytest
import torch

import fly_bproject_b200 as fb
from oracle import fly_oracle as fo
from tests.helpers import ATOL_ENV, ENV_CASES, RTOL_ENV, assert_obs_close, load_golden, oracle_tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_env(n, asset, full, root, dof, force, walking=False):
    src = fb.TensorStateSource(root.to(DEV).contiguous(), dof.to(DEV).contiguous(), force.to(DEV).contiguous())
    args = fb.make_args(num_envs=n, sim_device=DEV, asset=asset, full_actions=full, walking=walking)
    return fb.Fly(args, state_source=src)


@pytest.mark.parametrize("name,asset,n,full", ENV_CASES)
def test_env_kernels_match_reference_golden(name, asset, n, full):
    g = load_golden(name)
    env = make_env(n, asset, full, g["root"], g["dof_states"], g["force"])
    # pre-physics with no env flagged: scaling + DOF-target scatter only (fly.py:626-657)
    env.reset_buf.zero_()
    env.pre_physics(g["actions_in"].to(DEV))
    torch.testing.assert_close(env._actions_scaled.cpu(), g["actions_scaled"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(env.dof_targets.cpu(), g["dof_targets"], rtol=1e-6, atol=1e-6)
    assert env.actions.shape == (n * env.num_act, 1)
    # post-physics with the fixture's previous mask / counters / potentials
    env.reset_buf.copy_(g["reset_in"])
    env.progress_buf.copy_(g["progress_in"])
    env.potentials.copy_(g["potentials_in"])
    env.post_physics()
    torch.cuda.synchronize()
    assert torch.equal(env.reset_buf.cpu(), g["reset"])                       # bit-exact
    assert torch.equal(env.progress_buf.cpu(), g["progress"])
    assert env.reset_buf.dtype == torch.int64 and env.progress_buf.dtype == torch.int64
    assert_obs_close(env.obs_buf.cpu(), g["obs"], env.num_act)
    torch.testing.assert_close(env.reward_buf.cpu(), g["reward"], rtol=RTOL_ENV, atol=ATOL_ENV)
    assert torch.equal(env.potentials.cpu(), g["potentials"])                 # bit-exact (env_math.cuh: sumsq3_rn)
    assert torch.equal(env.prev_potentials.cpu(), g["prev_potentials"])
    torch.testing.assert_close(env.up_vec.cpu(), g["up_vec"], rtol=RTOL_ENV, atol=ATOL_ENV)
    torch.testing.assert_close(env.heading_vec.cpu(), g["heading_vec"], rtol=RTOL_ENV, atol=ATOL_ENV)
    # compacted ids == reset_buf.nonzero() (fly.py:447), ascending, int32; count on the device
    want_ids = g["reset"].nonzero().squeeze(-1).to(torch.int32)
    cnt = int(env.reset_count.item())
    assert cnt == want_ids.numel()
    assert torch.equal(env.reset_ids[:cnt].cpu(), want_ids)
    torch.testing.assert_close(env.reward_sum.cpu(), g["reward"].sum().view(1), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("n,asset,full,walking", [(16384, "limited_18", True, False), (100003, "limited_18", True, True),
                                                   (4099, "full_42", True, False), (1000, "full_42", False, True),
                                                   (1, "limited_18", True, False), (127, "limited_18", True, False)])
def test_env_step_matches_oracle(n, asset, full, walking):
    t = oracle_tables(asset, n, full)
    st = fo.synth_state(t, seed=n + 5)
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = st["root"].clone(), st["dof_states"].clone(), st["force"].clone()
    s.progress_buf = st["progress"].clone()
    g = torch.Generator().manual_seed(n)
    s.reset_buf = (torch.rand(n, generator=g) < 0.25).long()           # some envs flagged from the previous step
    s.potentials = -torch.rand(n, generator=g) * 60000.0
    env = make_env(n, asset, full, st["root"], st["dof_states"], st["force"], walking)
    env.progress_buf.copy_(s.progress_buf)
    env.reset_buf.copy_(s.reset_buf)
    env.potentials.copy_(s.potentials)
    env.prev_potentials.copy_(s.prev_potentials)

    # oracle: Fly.step order (fly.py:624-681) with no physics in between
    scaled, targets = fo.scale_actions(t, st["actions"])
    s.actions = scaled
    ids = fo.apply_reset(t, s)
    fo.post_physics(t, s, walking=walking)

    env.pre_physics(st["actions"].to(DEV))
    # reset bookkeeping of the flagged envs (fly.py:446-480)
    assert int(env.reset_buf.sum().item()) == 0
    torch.testing.assert_close(env.state.root.cpu(), s.root, rtol=0, atol=0)
    torch.testing.assert_close(env.state.dof_state.cpu(), s.dof_states, rtol=1e-6, atol=1e-7)
    env.simulate()
    env.post_physics()
    torch.cuda.synchronize()
    assert torch.equal(env.reset_buf.cpu(), s.reset_buf)
    assert torch.equal(env.progress_buf.cpu(), s.progress_buf)
    assert_obs_close(env.obs_buf.cpu(), s.obs_buf, env.num_act)
    # the walking total holds 2*(pot - prev_pot), a difference of O(6e4) potentials: they are bit-exact (just below), so
    # the walking reward is held to the same bar as the standing one
    torch.testing.assert_close(env.reward_buf.cpu(), s.reward_buf, rtol=RTOL_ENV, atol=ATOL_ENV)
    assert torch.equal(env.potentials.cpu(), s.potentials)
    assert torch.equal(env.prev_potentials.cpu(), s.prev_potentials)
    torch.testing.assert_close(env.dof_targets.cpu(), targets, rtol=1e-6, atol=1e-6)
    want_ids = s.reset_buf.nonzero().squeeze(-1).to(torch.int32)
    cnt = int(env.reset_count.item())
    assert cnt == want_ids.numel()
    assert torch.equal(env.reset_ids[:cnt].cpu(), want_ids)
    assert torch.equal(env.reset_ids[:cnt].cpu(), torch.sort(env.reset_ids[:cnt].cpu()).values)
    assert ids.dtype == torch.int32


def test_compaction_is_stable_over_repeated_launches():
    """The look-back scratch re-arms itself: many launches, all-set / none-set / ragged masks."""
    n = 50000
    t = oracle_tables("limited_18", n, True)
    st = fo.synth_state(t, seed=3)
    env = make_env(n, "limited_18", True, st["root"], st["dof_states"], st["force"])
    env._actions_scaled.copy_(st["actions"])
    for it in range(12):
        env.reset_buf.zero_()
        if it % 3 == 1:
            env.reset_buf.fill_(1)                      # every env flagged (the reference's first step, fly.py:175)
        env.progress_buf.zero_()
        env.post_physics()
        want = env.reset_buf.nonzero().squeeze(-1).to(torch.int32)
        cnt = int(env.reset_count.item())
        assert cnt == want.numel()
        assert torch.equal(env.reset_ids[:cnt], want)
    # a healthy, upright fly never resets: count 0 and ids untouched
    root = torch.zeros(n, 13)
    root[:, 2], root[:, 6] = 2.0, 1.0
    env2 = make_env(n, "limited_18", True, root, st["dof_states"], torch.zeros_like(st["force"]))
    env2.reset_buf.zero_()
    env2._actions_scaled.copy_(((t.lower + t.upper) / 2).repeat(n, 1))      # mid-range commands: no limit cost
    env2.post_physics()
    assert int(env2.reset_count.item()) == 0
    assert int(env2.reset_buf.sum().item()) == 0
    # closed forms: identity quaternion => up_proj = heading_proj = 1, roll = pitch = yaw = 0
    o = env2.obs_buf[0].cpu()
    assert o[10] == 1.0 and o[11] == 1.0 and o[7] == 0.0 and o[8] == 0.0 and o[12 + 54] == 0.0
    assert float(env2.reward_buf[0]) == pytest.approx(0.5)     # alive only: up 0 (1.4 < h < 2.1), no legs down


def test_step_api_and_errors():
    env = fb.Fly(fb.make_args(num_envs=300, sim_device=DEV))
    assert env.reset_buf.dtype == torch.int64 and int(env.reset_buf.sum()) == 300      # init ones (fly.py:175)
    a = torch.zeros(300, 18, device=DEV)
    env.step(a)
    assert int(env.progress_buf.min()) == 1
    with pytest.raises(fb.FlybError):
        env.step(torch.zeros(300, 17, device=DEV))
    with pytest.raises(fb.FlybError):
        env.step(torch.zeros(300, 18))                  # CPU tensor: no fallback
    assert env.dof_pos.shape == (300, 18) and env.root_tensor.shape == (300, 13)
    assert env.force_tensor.shape == (300 * 93, 3)


def test_step_tail_rides_on_post_physics():
    """FlybStepTail: the per-step variance annealing (ppo.py:236-237) and the Philox draw counter are applied exactly
    once per launch by the last CTA, whatever the grid size; without a tail nothing else is touched."""
    from fly_bproject_b200 import _lib
    for n in (64, 1000, 40000):
        env = fb.Fly(fb.make_args(num_envs=n, sim_device=DEV, state_pool=1, seed=n))
        var = torch.full((18,), 0.2, device=DEV)
        var[3] = 0.0100049                                  # lands on the floor after one step
        counter = torch.zeros(1, dtype=torch.int64, device=DEV)
        tail = _lib.StepTail(var.data_ptr(), 18, 0.00001, 0.01, counter.data_ptr())
        want = var.cpu().clone()
        for it in range(5):
            env.post_physics(tail=tail)
            want = fo.anneal_var(want)
        env.post_physics()                                  # no tail: variance and counter stay
        torch.cuda.synchronize()
        assert int(counter.item()) == 5
        torch.testing.assert_close(var.cpu(), want, rtol=0, atol=0)
        assert float(var[3]) == pytest.approx(0.01)


def test_lowgrav_variant_matches_reference_golden():
    """args.variant = "lowgrav": Fly.step = scale -> simulate -> reset -> obs / reward (flyLowGrav.py:617-683) with
    energy_cost_scale = 1 (flyLowGrav.py:44), against outputs of the executed flyLowGrav.compute_fly_* functions."""
    g = load_golden("env_lowgrav_d18_a18_n193.npz")
    n = 193
    src = fb.TensorStateSource(g["root"].to(DEV).contiguous(), g["dof_states"].to(DEV).contiguous(),
                               g["force"].to(DEV).contiguous())
    env = fb.Fly(fb.make_args(num_envs=n, sim_device=DEV, variant="lowgrav"), state_source=src)
    assert env.cfg.energy_cost_scale == 1.0 and env.reset_after_simulate
    env.reset_buf.copy_(g["reset_in"])
    env.progress_buf.copy_(g["progress_in"])
    env.potentials.copy_(g["potentials_in"])
    env.prev_potentials.copy_(g["potentials_in"])
    env.step(g["actions_in"].to(DEV))
    torch.cuda.synchronize()
    torch.testing.assert_close(env.dof_targets.cpu(), g["dof_targets"], rtol=1e-6, atol=1e-6)
    assert torch.equal(env.state.root.cpu(), g["root_after"])
    torch.testing.assert_close(env.state.dof_state.cpu(), g["dof_states_after"], rtol=1e-6, atol=1e-7)
    assert torch.equal(env.reset_buf.cpu(), g["reset"])
    assert torch.equal(env.progress_buf.cpu(), g["progress"])
    assert_obs_close(env.obs_buf.cpu(), g["obs"], env.num_act)
    torch.testing.assert_close(env.reward_buf.cpu(), g["reward"], rtol=RTOL_ENV, atol=ATOL_ENV)
    assert torch.equal(env.potentials.cpu(), g["potentials"])
    assert torch.equal(env.prev_potentials.cpu(), g["prev_potentials"])
    want_ids = g["reset"].nonzero().squeeze(-1).to(torch.int32)
    cnt = int(env.reset_count.item())
    assert cnt == want_ids.numel() and torch.equal(env.reset_ids[:cnt].cpu(), want_ids)


def test_walking_total_matches_reference_golden():
    """args.walking: the walking total + leg term (configs[2]) against the reference's reward function run with its own
    commented-out total (fly.py:747-748) switched back on (tests/golden/make_golden.py --walking-only)."""
    g = load_golden("env_walking_d18_a18_n211.npz")
    n = 211
    env = make_env(n, "limited_18", True, g["root"], g["dof_states"], g["force"], walking=True)
    env.reset_buf.zero_()
    env.pre_physics(g["actions_in"].to(DEV))
    env.reset_buf.copy_(g["reset_in"])
    env.progress_buf.copy_(g["progress_in"])
    env.potentials.copy_(g["potentials_in"])
    env.post_physics()
    torch.cuda.synchronize()
    assert torch.equal(env.reset_buf.cpu(), g["reset"])
    assert_obs_close(env.obs_buf.cpu(), g["obs"], env.num_act)
    assert torch.equal(env.potentials.cpu(), g["potentials"])    # bit-exact, so 2 * (potential - previous) is exact too
    torch.testing.assert_close(env.reward_buf.cpu(), g["reward"], rtol=RTOL_ENV, atol=ATOL_ENV)
    live = g["reset"] == 0
    err = (env.reward_buf.cpu() - g["reward"])[live].abs()
    assert float(err.median()) <= 1e-5                           # typical env: no ulp flips at all


def test_reset_method_resets_flagged_envs():
    """Fly.reset() (fly.py:446-480) performs the reset and reports whether any env was flagged."""
    n = 500
    t = oracle_tables("limited_18", n, True)
    st = fo.synth_state(t, seed=77)
    env = make_env(n, "limited_18", True, st["root"], st["dof_states"], st["force"])
    s = fo.EnvState(t)
    s.root, s.dof_states = st["root"].clone(), st["dof_states"].clone()
    g = torch.Generator().manual_seed(9)
    s.reset_buf = (torch.rand(n, generator=g) < 0.3).long()
    s.progress_buf = st["progress"].clone()
    s.potentials = -torch.rand(n, generator=g) * 1000.0
    s.prev_potentials = s.potentials.clone()
    env.reset_buf.copy_(s.reset_buf); env.progress_buf.copy_(s.progress_buf)
    env.potentials.copy_(s.potentials); env.prev_potentials.copy_(s.prev_potentials)
    fo.apply_reset(t, s)
    assert env.reset() is True
    torch.cuda.synchronize()
    assert int(env.reset_buf.sum()) == 0 and env.reset() is False
    assert torch.equal(env.progress_buf.cpu(), s.progress_buf)
    assert torch.equal(env.state.root.cpu(), s.root)
    torch.testing.assert_close(env.state.dof_state.cpu(), s.dof_states, rtol=1e-6, atol=1e-7)
    assert torch.equal(env.potentials.cpu(), s.potentials)
    assert torch.equal(env.prev_potentials.cpu(), s.prev_potentials)
