"""Freeze golden vectors by EXECUTING THE REFERENCE'S OWN CODE on CPU torch.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
The reference has no tests or golden vectors of its own; these fixtures are outputs of its
unmodified ``fly.compute_fly_observations`` / ``fly.compute_fly_reward2`` (fly.py:685-805) and of its ``ppo.Net`` /
``ppo.PPO.run/make_data/update`` (ppo.py:10-264) driven through ``oracle/ref_harness.py`` on seeded synthetic inputs.
They travel to the GPU box, where /root/reference does not exist.
"""
import contextlib
import io
import json
import os
import sys
import types
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

from oracle import fly_oracle as fo  # noqa: E402
from oracle import ref_harness  # noqa: E402

CKPT = os.path.join(ref_harness.REFERENCE_ROOT, "saves/save7_1_23/save1_standing_18dofs_7800.pth")
TABLES = json.load(open(os.path.join(ROOT, "fly_bproject_b200/data/nmf_tables.json")))["assets"]


def tables_for(asset: str, n: int, full_actions: bool) -> fo.EnvTables:
    t = TABLES[asset]
    aidx = list(range(t["num_dof"])) if full_actions else t["action_index_18"]
    return fo.EnvTables(n, t["lower"], t["upper"], t["initial_pos"], aidx, t["abdomen_local"], t["leg_tip_local"],
                        t["num_bodies"])


def env_case(fly, asset: str, n: int, full_actions: bool, seed: int, prev_reset_frac: float, reward_fn=None):
    """`reward_fn`: stands in for fly.compute_fly_reward2 — the walking total (ref_harness.load_walking_reward)."""
    t = tables_for(asset, n, full_actions)
    st = fo.synth_state(t, seed)
    g = torch.Generator().manual_seed(seed + 99)
    reset_prev = (torch.rand(n, generator=g) < prev_reset_frac).long()
    potentials = -torch.rand(n, generator=g) * 60000.0
    scaled, targets = fo.scale_actions(t, st["actions"])
    dof = st["dof_states"].view(n, t.num_dof, 2)
    a, o = t.num_act, t.num_obs
    if reward_fn is not None:
        # the walking total holds 2 * (potential - previous potential): previous potentials one plausible step behind the
        # ones this state yields (a first evaluation tells them), so that the term is O(1) as it is in a live rollout
        pot0 = fly.compute_fly_observations(
            torch.zeros(n, o), st["root"], t.targets, potentials.clone(), t.inv_start_rot, dof[..., 0], dof[..., 1],
            t.lower, t.upper, 0.2, scaled.view(n, -1), 1 / 60, t.basis0, t.basis1, 2, t.action_idx_one, st["force"],
            t.leg_tip_ids, n)[1]
        potentials = (pot0 - (torch.rand(n, generator=g) * 3.0 - 1.0)).clone()
    obs, pot, prev, up, head = fly.compute_fly_observations(
        torch.zeros(n, o), st["root"], t.targets, potentials.clone(), t.inv_start_rot, dof[..., 0], dof[..., 1],
        t.lower, t.upper, 0.2, scaled.view(n, -1), 1 / 60, t.basis0, t.basis1, 2, t.action_idx_one, st["force"],
        t.leg_tip_ids, n)
    progress = st["progress"].clone() + 1                                   # fly.py:678
    progress[progress == 0] = 1                                             # fly.py:415-416
    rew, reset = (reward_fn or fly.compute_fly_reward2)(
        obs, reset_prev.clone(), progress, scaled.view(n, -1), 0.75, 0.5, pot, prev, 0.005, 0.005, 0.1, 1.1, 6, -2.0,
        1500, t.action_idx_one, t.upper, t.lower, n, st["root"][:, 3:7], a, st["force"], t.abdomen_ids,
        t.leg_tip_ids)
    return {
        "root": st["root"], "dof_states": st["dof_states"], "force": st["force"], "progress_in": st["progress"],
        "reset_in": reset_prev, "potentials_in": potentials, "actions_in": st["actions"],
        "actions_scaled": scaled.view(n, a), "dof_targets": targets,
        "obs": obs, "potentials": pot, "prev_potentials": prev, "up_vec": up, "heading_vec": head,
        "progress": progress, "reward": rew, "reset": reset,
    }


def lowgrav_case(mod, asset: str, n: int, full_actions: bool, seed: int, prev_reset_frac: float):
    """One `Fly.step` of the low-gravity variant (flyLowGrav.py:617-683) on a synthetic post-simulate state: scale
    actions, [simulate], reset() of the flagged envs (oracle restatement of flyLowGrav.py:446-480 — the method itself
    needs gymapi), then flyLowGrav.compute_fly_observations / compute_fly_reward2 EXECUTED from the reference with
    energy_cost_scale = 1 (flyLowGrav.py:44): the observation sees the reset rows."""
    t = tables_for(asset, n, full_actions)
    st = fo.synth_state(t, seed)
    g = torch.Generator().manual_seed(seed + 99)
    reset_prev = (torch.rand(n, generator=g) < prev_reset_frac).long()
    potentials = -torch.rand(n, generator=g) * 60000.0
    s = fo.EnvState(t)
    s.root, s.dof_states, s.force = st["root"].clone(), st["dof_states"].clone(), st["force"].clone()
    s.progress_buf, s.reset_buf, s.potentials = st["progress"].clone(), reset_prev.clone(), potentials.clone()
    s.prev_potentials = potentials.clone()
    scaled, targets = fo.scale_actions(t, st["actions"])
    ids = fo.apply_reset(t, s)                                              # after simulate (flyLowGrav.py:661-663)
    dof = s.dof_states.view(n, t.num_dof, 2)
    a, o = t.num_act, t.num_obs
    obs, pot, prev, up, head = mod.compute_fly_observations(
        torch.zeros(n, o), s.root, t.targets, s.potentials.clone(), t.inv_start_rot, dof[..., 0], dof[..., 1],
        t.lower, t.upper, 0.2, scaled.view(n, -1), 1 / 60, t.basis0, t.basis1, 2, t.action_idx_one, s.force,
        t.leg_tip_ids, n)
    progress = s.progress_buf.clone() + 1
    progress[progress == 0] = 1
    rew, reset = mod.compute_fly_reward2(
        obs, s.reset_buf.clone(), progress, scaled.view(n, -1), 0.75, 0.5, pot, prev, 0.005, 1, 0.1, 1.1, 6, -2.0,
        1500, t.action_idx_one, t.upper, t.lower, n, s.root[:, 3:7], a, s.force, t.abdomen_ids, t.leg_tip_ids)
    return {
        "root": st["root"], "dof_states": st["dof_states"], "force": st["force"], "progress_in": st["progress"],
        "reset_in": reset_prev, "potentials_in": potentials, "actions_in": st["actions"],
        "actions_scaled": scaled.view(n, a), "dof_targets": targets, "reset_ids": ids,
        "root_after": s.root, "dof_states_after": s.dof_states,
        "obs": obs, "potentials": pot, "prev_potentials": prev, "up_vec": up, "heading_vec": head,
        "progress": progress, "reward": rew, "reset": reset,
    }


def save(name, d):
    arrs = {k: (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in d.items()}
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrs)
    print(f"{name}: {os.path.getsize(path) / 1024:.0f} KiB  keys={len(arrs)}")


def ppo_case(n=64, t_roll=32, mc=2):
    o, a = 73, 18

    class SynthEnv:
        def __init__(self, args):
            self.num_act, self.num_obs, self.end = a, o, False
            self.g = torch.Generator().manual_seed(7)
            self.obs_buf = torch.randn(n, o, generator=self.g)
            self.reward_buf = torch.randn(n, generator=self.g)
            self.reset_buf = torch.ones(n, dtype=torch.long)

        def step(self, act):
            self.obs_buf = torch.randn(n, o, generator=self.g)
            self.reward_buf = torch.randn(n, generator=self.g)
            self.reset_buf = (torch.rand(n, generator=self.g) < 0.3).long()

    ppo = ref_harness.load_ppo(SynthEnv)
    args = types.SimpleNamespace(num_envs=n, sim_device="cpu", load=False, load_path=None, save=False,
                                 save_path=None, testing=False)
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        agent = ppo.PPO(args)
    agent.net.load_state_dict(torch.load(CKPT, map_location="cpu"))
    agent.optim = torch.optim.Adam(agent.net.parameters(), lr=agent.lr)
    agent.mini_chunk_size, agent.rollout_size = mc, t_roll
    for name in ["all_obs", "all_acts", "all_next_obs", "all_reward", "all_done", "all_log_prob", "all_advantage"]:
        setattr(agent, name, getattr(agent, name)[:t_roll].clone())
    out = {}
    params0 = {k: v.clone() for k, v in agent.net.state_dict().items()}
    for k, v in params0.items():
        out["w0/" + k] = v

    # --- Net.pi / Net.v on a seeded batch (ppo.py:88-102)
    x = torch.randn(257, o, generator=torch.Generator().manual_seed(11)) * 1.5
    with torch.no_grad():
        out["net_x"], out["net_pi"], out["net_v"] = x, agent.net.pi(x), agent.net.v(x)

    # --- one act step with known noise (ppo.py:210-220): MVN.sample() == mu + sqrt(var)*randn under the same state
    torch.manual_seed(123)
    rng = torch.get_rng_state()
    out["act_noise"] = torch.randn(n, a)
    torch.set_rng_state(rng)
    out["act_obs"] = agent.env.obs_buf.clone()
    out["act_var"] = agent.action_var.clone()
    with contextlib.redirect_stdout(sink):
        agent.run()
    out["act_action"], out["act_logp"] = agent.all_acts[0].clone(), agent.all_log_prob[0].clone()

    # --- fill the rollout, then capture make_data / update (ppo.py:157-202)
    with contextlib.redirect_stdout(sink):
        for _ in range(t_roll - 2):
            agent.run()
    cap = {}
    real_make_data = agent.make_data

    def spy_make_data():
        res = real_make_data()
        if "target" not in cap:
            cap.update(obs=agent.all_obs.clone(), acts=agent.all_acts.clone(), next=agent.all_next_obs.clone(),
                       rew=agent.all_reward.clone(), done=agent.all_done.clone(), logp=agent.all_log_prob.clone(),
                       var=agent.action_var.clone(), target=res[3].clone(), adv=res[4].clone())
        return res

    agent.make_data = spy_make_data
    snaps = {}
    real_step = agent.optim.step

    def spy_step(*a_, **k_):
        # grads after clip_grad_norm_, before Adam (ppo.py:198-199)
        i = agent.optim_step
        if i in (0,):
            for name, prm in agent.net.named_parameters():
                snaps[f"g{i}/" + name] = prm.grad.clone()
        r = real_step(*a_, **k_)
        if i in (0, 2, 74):
            for name, prm in agent.net.named_parameters():
                snaps[f"w{i + 1}/" + name] = prm.detach().clone()
                state = agent.optim.state[prm]
                snaps[f"m{i + 1}/" + name] = state["exp_avg"].clone()
                snaps[f"v{i + 1}/" + name] = state["exp_avg_sq"].clone()
        return r

    agent.optim.step = spy_step
    with contextlib.redirect_stdout(sink):
        agent.run()
    assert agent.optim_step == 75, agent.optim_step                        # log.txt:49: 5 epochs x 15 minibatches
    assert tuple(cap["done"].shape) == (n, 1) and cap["done"].dtype == torch.int64   # ppo.py:230 quirk
    out.update({"roll_" + k: v for k, v in cap.items()})
    out.update(snaps)
    # loss of the very first minibatch, recomputed with the reference net restored to w0 (ppo.py:184-197)
    net0 = ppo.Net(o, a)
    net0.load_state_dict(params0)
    from torch.distributions import MultivariateNormal
    mu = net0.pi(cap["obs"][0:mc])
    dist = MultivariateNormal(mu, scale_tril=torch.linalg.cholesky(torch.diag(cap["var"])))
    ratio = torch.exp(dist.log_prob(cap["acts"][0:mc]) - cap["logp"][0:mc]).unsqueeze(-1)
    adv = cap["adv"][0:mc]
    loss = -torch.min(ratio * adv, torch.clamp(ratio, 0.8, 1.2) * adv) + \
        torch.nn.functional.smooth_l1_loss(net0.v(cap["obs"][0:mc]), cap["target"][0:mc])
    out["loss0"] = loss.mean().detach()
    out["meta"] = np.array([n, t_roll, mc])
    return out


def main():
    assert ref_harness.available(), "reference checkout missing"
    torch.manual_seed(0)
    fly = ref_harness.load_fly()
    if "--lowgrav-only" in sys.argv:        # later addition: leaves the round-1 fixtures byte-identical
        save("env_lowgrav_d18_a18_n193.npz",
             lowgrav_case(ref_harness.load_fly_lowgrav(), "limited_18", 193, True, 2024, 0.25))
        return
    if "--walking-only" in sys.argv:        # later addition, same rule
        save("env_walking_d18_a18_n211.npz",
             env_case(fly, "limited_18", 211, True, 515, 0.2, reward_fn=ref_harness.load_walking_reward()))
        return
    if "--walking-literal-only" in sys.argv:   # later addition, same rule: the literal un-commenting (no leg term); same
        lit = env_case(fly, "limited_18", 211, True, 515, 0.2,   # inputs as env_walking_d18_a18_n211.npz, outputs only
                       reward_fn=ref_harness.load_walking_reward(leg_term=False))
        save("env_walking_literal_d18_a18_n211.npz", {k: lit[k] for k in ("reward", "reset", "potentials_in")})
        return
    save("env_lowgrav_d18_a18_n193.npz", lowgrav_case(ref_harness.load_fly_lowgrav(), "limited_18", 193, True, 2024, 0.25))
    save("env_walking_d18_a18_n211.npz",
         env_case(fly, "limited_18", 211, True, 515, 0.2, reward_fn=ref_harness.load_walking_reward()))
    save("env_d18_a18_n257.npz", env_case(fly, "limited_18", 257, True, 1234, 0.2))
    save("env_d42_a18_n131.npz", env_case(fly, "full_42", 131, False, 4321, 0.1))
    save("env_d42_a42_n131.npz", env_case(fly, "full_42", 131, True, 999, 0.0))
    save("ppo_n64_t32.npz", ppo_case())


if __name__ == "__main__":
    main()
