"""CPU: the travelling oracle (oracle/fly_oracle.py) against the golden vectors frozen from the EXECUTED reference
(tests/golden/make_golden.py ran reference fly.py:685-805 and ppo.py:10-264 unmodified)."""
import pytest
import torch

from oracle import fly_oracle as fo
from tests.helpers import ENV_CASES, load_golden, oracle_tables


@pytest.mark.parametrize("name,asset,n,full", ENV_CASES)
def test_env_oracle_matches_reference_golden(name, asset, n, full):
    g = load_golden(name)
    t = oracle_tables(asset, n, full)
    scaled, targets = fo.scale_actions(t, g["actions_in"])
    assert torch.equal(scaled.view(n, -1), g["actions_scaled"])
    assert torch.equal(targets, g["dof_targets"])
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = g["root"].clone(), g["dof_states"].clone(), g["force"].clone()
    s.progress_buf, s.reset_buf = g["progress_in"].clone(), g["reset_in"].clone()
    s.potentials = g["potentials_in"].clone()
    s.actions = scaled
    fo.post_physics(t, s)
    # same machine class, same op order -> bit-identical
    for key, got in [("obs", s.obs_buf), ("potentials", s.potentials), ("prev_potentials", s.prev_potentials),
                     ("up_vec", s.up_vec), ("heading_vec", s.heading_vec), ("reward", s.reward_buf)]:
        torch.testing.assert_close(got, g[key], rtol=1e-6, atol=1e-6, msg=key)
    assert torch.equal(s.reset_buf, g["reset"])
    assert torch.equal(s.progress_buf, g["progress"])
    # every termination branch is exercised by the fixture
    assert 0 < int(g["reset"].sum()) < n


def test_walking_total_matches_reference_golden():
    """configs[2]'s reward against the fixture frozen from the reference's reward function with its commented-out walking
    total (fly.py:747-748) switched back on (oracle/ref_harness.py:load_walking_reward)."""
    g = load_golden("env_walking_d18_a18_n211.npz")
    n = 211
    t = oracle_tables("limited_18", n, True)
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = g["root"].clone(), g["dof_states"].clone(), g["force"].clone()
    s.progress_buf, s.reset_buf = g["progress_in"].clone(), g["reset_in"].clone()
    s.potentials = g["potentials_in"].clone()
    s.actions, _ = fo.scale_actions(t, g["actions_in"])
    fo.post_physics(t, s, walking=True)
    torch.testing.assert_close(s.obs_buf, g["obs"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(s.reward_buf, g["reward"], rtol=1e-6, atol=1e-6)
    assert torch.equal(s.reset_buf, g["reset"]) and 0 < int(g["reset"].sum()) < n
    live = g["reset"] == 0
    progress2 = 2 * (g["potentials"] - g["prev_potentials"])[live]
    assert 0.5 < float(progress2.abs().max()) < 8.0          # the term is there, at a live rollout's scale


def test_walking_total_literal_reading_matches_reference_golden():
    """The same inputs, the LITERAL un-commenting of fly.py:747-748 (the nested `#+ leg_ground_reward` stays a comment):
    fixture env_walking_literal_d18_a18_n211.npz against the oracle with `walking_leg_term=False`.  Keeps both readings
    of the dead line pinned (the default one, with the leg term, is an interpretation: oracle/ref_harness.py)."""
    g, lit = load_golden("env_walking_d18_a18_n211.npz"), load_golden("env_walking_literal_d18_a18_n211.npz")
    assert torch.equal(lit["potentials_in"], g["potentials_in"])        # same inputs as the other fixture
    n = 211
    t = oracle_tables("limited_18", n, True)
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = g["root"].clone(), g["dof_states"].clone(), g["force"].clone()
    s.progress_buf, s.reset_buf = g["progress_in"].clone(), g["reset_in"].clone()
    s.potentials = g["potentials_in"].clone()
    s.actions, _ = fo.scale_actions(t, g["actions_in"])
    fo.post_physics(t, s, walking=True, walking_leg_term=False)
    torch.testing.assert_close(s.reward_buf, lit["reward"], rtol=1e-6, atol=1e-6)
    assert torch.equal(s.reset_buf, lit["reset"]) and torch.equal(lit["reset"], g["reset"])
    assert not torch.equal(lit["reward"], g["reward"])                   # the leg term is what separates them


def _params(g, tag="w0"):
    return {k: g[f"{tag}/{k}"].clone() for k in fo.PARAM_ORDER}


def test_net_and_act_match_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    p = _params(g)
    torch.testing.assert_close(fo.pi(p, g["net_x"]), g["net_pi"], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(fo.v(p, g["net_x"]), g["net_v"], rtol=1e-5, atol=1e-5)
    a, logp, _ = fo.act(p, g["act_obs"], g["act_var"], g["act_noise"])
    torch.testing.assert_close(a, g["act_action"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(logp, g["act_logp"], rtol=1e-5, atol=1e-5)


def test_gae_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    p = _params(g)
    target, adv = fo.make_data(p, g["roll_obs"], g["roll_next"], g["roll_rew"], g["roll_done"])
    torch.testing.assert_close(target, g["roll_target"], rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(adv, g["roll_adv"], rtol=1e-5, atol=1e-5)
    # the ppo.py:230 quirk: done is the LAST step's (N,1) mask
    assert tuple(g["roll_done"].shape) == (64, 1)
    # all_next_obs[t] is all_obs[t+1] (ppo.py:210, 228): what the (T+1)-row value forward relies on
    assert torch.equal(g["roll_next"][:-1], g["roll_obs"][1:])


def test_update_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    p = _params(g)
    st = fo.AdamState(p)
    target, adv = g["roll_target"], g["roll_adv"]
    loss, _, grads = fo.minibatch_step(p, st, g["roll_obs"][0:mc], g["roll_acts"][0:mc], g["roll_logp"][0:mc],
                                       target[0:mc], adv[0:mc], g["roll_var"])
    torch.testing.assert_close(torch.tensor(loss), g["loss0"], rtol=1e-5, atol=1e-6)
    total = torch.sqrt(sum((x.double() ** 2).sum() for x in grads.values()))
    coef = min(1.0, 1.0 / (float(total) + 1e-6))
    for k in fo.PARAM_ORDER:
        torch.testing.assert_close(grads[k] * coef, g["g0/" + k], rtol=1e-4, atol=1e-7, msg="grad " + k)
        torch.testing.assert_close(p[k], g["w1/" + k], rtol=1e-5, atol=1e-6, msg="param " + k)
        torch.testing.assert_close(st.m[k], g["m1/" + k], rtol=1e-4, atol=1e-8, msg="exp_avg " + k)
        torch.testing.assert_close(st.v[k], g["v1/" + k], rtol=1e-4, atol=1e-12, msg="exp_avg_sq " + k)


def test_three_steps_match_reference_golden():
    """w3 / m3 / v3: the reference's parameters and Adam moments after optimiser steps 1-3 (ppo.py:178-202)."""
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    p = _params(g)
    st = fo.AdamState(p)
    for i in range(3):
        k0, k1 = i * mc, (i + 1) * mc
        fo.minibatch_step(p, st, g["roll_obs"][k0:k1], g["roll_acts"][k0:k1], g["roll_logp"][k0:k1],
                          g["roll_target"][k0:k1], g["roll_adv"][k0:k1], g["roll_var"])
    for k in fo.PARAM_ORDER:
        torch.testing.assert_close(p[k], g["w3/" + k], rtol=1e-5, atol=1e-6, msg="param " + k)
        torch.testing.assert_close(st.m[k], g["m3/" + k], rtol=1e-4, atol=1e-8, msg="exp_avg " + k)
        torch.testing.assert_close(st.v[k], g["v3/" + k], rtol=1e-4, atol=1e-12, msg="exp_avg_sq " + k)


def test_full_update_75_steps_matches_reference_golden():
    g = load_golden("ppo_n64_t32.npz")
    n, t_roll, mc = [int(x) for x in g["meta"]]
    p = _params(g)
    st = fo.AdamState(p)
    losses = fo.update(p, st, g["roll_obs"], g["roll_acts"], g["roll_logp"], g["roll_next"], g["roll_rew"],
                       g["roll_done"], g["roll_var"], mc)
    assert len(losses) == 75 and st.step == 75           # log.txt:49: 5 epochs x 15 minibatches
    for k in fo.PARAM_ORDER:
        torch.testing.assert_close(p[k], g["w75/" + k], rtol=2e-3, atol=2e-4, msg="param " + k)


def test_lowgrav_step_matches_reference_golden():
    """Low-gravity variant: reset AFTER simulate (flyLowGrav.py:657-663), energy_cost_scale = 1 (flyLowGrav.py:44);
    the fixture's obs / reward are outputs of flyLowGrav.compute_fly_* executed from the reference."""
    g = load_golden("env_lowgrav_d18_a18_n193.npz")
    n = 193
    t = oracle_tables("limited_18", n, True)
    s = fo.EnvState(t)
    s.progress_buf, s.reset_buf = g["progress_in"].clone(), g["reset_in"].clone()
    s.potentials, s.prev_potentials = g["potentials_in"].clone(), g["potentials_in"].clone()

    def simulate(state):
        state.root, state.dof_states, state.force = g["root"].clone(), g["dof_states"].clone(), g["force"].clone()

    targets, ids = fo.env_step(t, s, g["actions_in"], simulate, variant="lowgrav")
    assert torch.equal(targets, g["dof_targets"]) and torch.equal(ids, g["reset_ids"])
    assert torch.equal(s.root, g["root_after"]) and torch.equal(s.dof_states, g["dof_states_after"])
    for key, got in [("obs", s.obs_buf), ("potentials", s.potentials), ("prev_potentials", s.prev_potentials),
                     ("reward", s.reward_buf), ("up_vec", s.up_vec), ("heading_vec", s.heading_vec)]:
        torch.testing.assert_close(got, g[key], rtol=1e-6, atol=1e-6, msg=key)
    assert torch.equal(s.reset_buf, g["reset"]) and torch.equal(s.progress_buf, g["progress"])
    # the envs reset inside this step start their episode here: progress 1, origin height 2.0 in the observation
    flagged = g["reset_in"].bool()
    assert int(flagged.sum()) > 10
    assert torch.equal(g["progress"][flagged], torch.ones_like(g["progress"][flagged]))
    assert torch.equal(g["obs"][flagged, 0], torch.full((int(flagged.sum()),), 2.0))
    # standard order on the same inputs differs exactly on those envs (reset before simulate: the obs sees the raw state)
    s2 = fo.EnvState(t)
    s2.progress_buf, s2.reset_buf = g["progress_in"].clone(), g["reset_in"].clone()
    s2.potentials, s2.prev_potentials = g["potentials_in"].clone(), g["potentials_in"].clone()
    fo.env_step(t, s2, g["actions_in"], simulate, variant="standard")
    assert torch.equal(s2.obs_buf[~flagged], s.obs_buf[~flagged])
    assert not torch.equal(s2.obs_buf[flagged], s.obs_buf[flagged])
