"""CPU: the travelling oracle against the reference's OWN code, executed here.

Runs wherever the reference's files exist — `/root/reference` in the build container, or the byte-for-byte staged
copy `oracle/_ref` (oracle/make_ref.py) that travels with the gpurun snapshot — and is skipped elsewhere (the committed
fixtures under tests/golden/ then carry the pin: tests/test_oracle_golden.py).  On fresh seeded inputs (not the ones
the fixtures were frozen from) the oracle's restatement must reproduce `fly.compute_fly_observations`,
`fly.compute_fly_reward2` (reference fly.py:685-805), `flyLowGrav`'s twins, `ppo.Net.pi / v` (ppo.py:88-102) and
`ppo.PPO.make_data` (ppo.py:157-171) — same op order on the same machine, so masks are bit-exact and fp32 values agree
to 1e-6.
"""
import contextlib
import io
import types

import pytest
import torch

from oracle import fly_oracle as fo
from oracle import ref_harness
from tests.helpers import oracle_tables

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="reference files not present (oracle/make_ref.py)")


def _run_reference_env(mod, t, st, reset_prev, potentials, energy_cost_scale=0.005):
    n, a, o = t.num_envs, t.num_act, t.num_obs
    scaled, _ = fo.scale_actions(t, st["actions"])
    dof = st["dof_states"].view(n, t.num_dof, 2)
    obs, pot, prev, up, head = mod.compute_fly_observations(
        torch.zeros(n, o), st["root"], t.targets, potentials.clone(), t.inv_start_rot, dof[..., 0], dof[..., 1],
        t.lower, t.upper, 0.2, scaled.view(n, -1), 1 / 60, t.basis0, t.basis1, 2, t.action_idx_one, st["force"],
        t.leg_tip_ids, n)
    progress = st["progress"].clone() + 1
    progress[progress == 0] = 1
    rew, reset = mod.compute_fly_reward2(
        obs, reset_prev.clone(), progress, scaled.view(n, -1), 0.75, 0.5, pot, prev, 0.005, energy_cost_scale, 0.1,
        1.1, 6, -2.0, 1500, t.action_idx_one, t.upper, t.lower, n, st["root"][:, 3:7], a, st["force"], t.abdomen_ids,
        t.leg_tip_ids)
    return scaled, obs, pot, prev, up, head, progress, rew, reset


@pytest.mark.parametrize("asset,n,full,seed", [("limited_18", 513, True, 31), ("full_42", 97, False, 32),
                                               ("full_42", 97, True, 33)])
@pytest.mark.parametrize("variant", ["fly", "flyLowGrav"])
def test_env_oracle_equals_executed_reference(asset, n, full, seed, variant):
    mod = ref_harness.load_fly() if variant == "fly" else ref_harness.load_fly_lowgrav()
    ecs = 0.005 if variant == "fly" else 1.0                  # flyLowGrav.py:44
    t = oracle_tables(asset, n, full)
    st = fo.synth_state(t, seed)
    g = torch.Generator().manual_seed(seed + 1)
    reset_prev = (torch.rand(n, generator=g) < 0.15).long()
    potentials = -torch.rand(n, generator=g) * 60000.0
    scaled, obs, pot, prev, up, head, progress, rew, reset = _run_reference_env(mod, t, st, reset_prev, potentials, ecs)

    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = st["root"].clone(), st["dof_states"].clone(), st["force"].clone()
    s.progress_buf, s.reset_buf, s.potentials = st["progress"].clone(), reset_prev.clone(), potentials.clone()
    s.actions = scaled
    fo.post_physics(t, s, energy_cost_scale=ecs)
    for name, got, want in [("obs", s.obs_buf, obs), ("potentials", s.potentials, pot),
                            ("prev_potentials", s.prev_potentials, prev), ("up_vec", s.up_vec, up),
                            ("heading_vec", s.heading_vec, head), ("reward", s.reward_buf, rew)]:
        torch.testing.assert_close(got, want, rtol=1e-6, atol=1e-6, msg=name)
    assert torch.equal(s.reset_buf, reset)
    assert torch.equal(s.progress_buf, progress)
    assert 0 < int(reset.sum()) < n


@pytest.mark.parametrize("leg_term", [True, False])
@pytest.mark.parametrize("asset,n,full,seed", [("limited_18", 513, True, 41), ("full_42", 97, True, 42)])
def test_walking_total_equals_the_references_commented_out_line_executed(asset, n, full, seed, leg_term):
    """configs[2]'s reward: the oracle's `walking=True` branch against compute_fly_reward2 with the reference's own
    commented-out total (fly.py:747-748) switched back on (ref_harness.load_walking_reward) — in both readings: with the
    nested `#+ leg_ground_reward` re-enabled as well (an interpretation, the oracle's and the kernels' default) and as
    the literal un-commenting, which leaves the leg term out."""
    mod = ref_harness.load_fly()
    walking = ref_harness.load_walking_reward(leg_term=leg_term)
    t = oracle_tables(asset, n, full)
    st = fo.synth_state(t, seed)
    g = torch.Generator().manual_seed(seed + 1)
    reset_prev = (torch.rand(n, generator=g) < 0.15).long()
    potentials = -torch.rand(n, generator=g) * 60000.0
    a = t.num_act
    scaled, obs, pot, prev, up, head, progress, rew_standing, reset_standing = _run_reference_env(
        mod, t, st, reset_prev, potentials)
    rew, reset = walking(
        obs, reset_prev.clone(), progress, scaled.view(n, -1), 0.75, 0.5, pot, prev, 0.005, 0.005, 0.1, 1.1, 6, -2.0,
        1500, t.action_idx_one, t.upper, t.lower, n, st["root"][:, 3:7], a, st["force"], t.abdomen_ids, t.leg_tip_ids)
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = st["root"].clone(), st["dof_states"].clone(), st["force"].clone()
    s.progress_buf, s.reset_buf, s.potentials = st["progress"].clone(), reset_prev.clone(), potentials.clone()
    s.actions = scaled
    fo.post_physics(t, s, walking=True, walking_leg_term=leg_term)
    assert torch.equal(s.reset_buf, reset) and torch.equal(reset, reset_standing)
    # same operations in the same order on the same obs: the totals agree to the last bit or one rounding
    torch.testing.assert_close(s.reward_buf, rew, rtol=1e-6, atol=1e-6)
    live = reset == 0
    assert float((rew - rew_standing)[live].abs().max()) > 1.0      # it IS a different total (progress term)
    # the two readings differ by exactly the leg-contact term (0.1 per leg tip on the ground, fly.py:744) on live envs
    other = ref_harness.load_walking_reward(leg_term=not leg_term)(
        obs, reset_prev.clone(), progress, scaled.view(n, -1), 0.75, 0.5, pot, prev, 0.005, 0.005, 0.1, 1.1, 6, -2.0,
        1500, t.action_idx_one, t.upper, t.lower, n, st["root"][:, 3:7], a, st["force"], t.abdomen_ids, t.leg_tip_ids)[0]
    legs = obs[:, 13 + 3 * a:19 + 3 * a].sum(dim=1) * 0.1
    alive = rew != -2.0
    sign = 1.0 if leg_term else -1.0
    # (this test's previous potentials are random, so the totals are O(1e5): an ulp there is 0.0078)
    torch.testing.assert_close((rew - other)[alive], sign * legs[alive], rtol=0, atol=0.02)
    assert float(legs[alive].max()) > 0.0


def _reference_agent(n, o, a):
    class SynthEnv:
        def __init__(self, args):
            self.num_act, self.num_obs, self.end = a, o, False
            self.obs_buf = torch.zeros(n, o)
            self.reward_buf = torch.zeros(n)
            self.reset_buf = torch.ones(n, dtype=torch.long)

        def step(self, act):
            pass

    ppo = ref_harness.load_ppo(SynthEnv)
    args = types.SimpleNamespace(num_envs=n, sim_device="cpu", load=False, load_path=None, save=False, save_path=None,
                                 testing=False)
    with contextlib.redirect_stdout(io.StringIO()):
        return ppo.PPO(args)


def test_net_and_make_data_equal_executed_reference():
    n, o, a, t_roll = 48, 73, 18, 12
    torch.manual_seed(5)
    agent = _reference_agent(n, o, a)
    p = {k: v.detach().clone() for k, v in agent.net.state_dict().items()}
    g = torch.Generator().manual_seed(6)
    x = torch.randn(301, o, generator=g) * 1.5
    with torch.no_grad():
        torch.testing.assert_close(fo.pi(p, x), agent.net.pi(x), rtol=1e-6, atol=1e-6)
        torch.testing.assert_close(fo.v(p, x), agent.net.v(x), rtol=1e-6, atol=1e-6)
    # make_data on a seeded buffer, with the ppo.py:230 quirk ((N,1) done mask of the last step)
    agent.all_obs = torch.randn(t_roll, n, o, generator=g)
    agent.all_next_obs = torch.cat([agent.all_obs[1:], torch.randn(1, n, o, generator=g)], 0)
    agent.all_reward = torch.randn(t_roll, n, 1, generator=g)
    agent.all_done = (torch.rand(n, 1, generator=g) < 0.7).long()
    agent.all_acts = torch.randn(t_roll, n, a, generator=g).clip(-1, 1)
    agent.all_log_prob = torch.randn(t_roll, n, generator=g)
    agent.all_advantage = torch.zeros(t_roll, n, 1)
    agent.rollout_size = t_roll
    res = agent.make_data()
    target, adv = fo.make_data(p, agent.all_obs, agent.all_next_obs, agent.all_reward, agent.all_done)
    torch.testing.assert_close(target, res[3], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(adv, res[4], rtol=1e-5, atol=1e-5)


def test_staged_copy_is_byte_identical_to_the_checkout():
    """When both exist (build container), oracle/_ref must be the unmodified reference (sha256 per file)."""
    import json
    import os
    from oracle import make_ref
    man = os.path.join(make_ref.DST, "MANIFEST.json")
    if not (os.path.isfile(man) and os.path.isfile(os.path.join(make_ref.SRC, "fly.py"))):
        pytest.skip("needs both /root/reference and oracle/_ref")
    files = json.load(open(man))["files"]
    assert set(files) == set(make_ref.FILES)
    for rel, digest in files.items():
        assert make_ref.sha256(os.path.join(make_ref.SRC, rel)) == digest == make_ref.sha256(os.path.join(make_ref.DST, rel))
