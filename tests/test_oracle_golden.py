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
