"""GPU: the drop-in boundary exercised the way INTEGRATION.md describes it — by the REFERENCE's own code.

(1) The reference's `ppo.PPO` (its net, its sampling, its buffers: reference ppo.py:105-264, unmodified) drives
    `fly_bproject_b200.Fly` as its env ("keep the reference's ppo.py, swap only Fly"); what the env hands back after
    every `run()` is checked against the oracle fed with the actions the reference agent chose.
(2) The reference's `trainer.py` (trainer.py:1-49, unmodified, executed with runpy) drives `fly_bproject_b200.PPO`
    through one whole rollout + update (the one-line change: `from ppo import PPO` resolves to the drop-in).
The reference files are the staged copies under oracle/_ref (or /root/reference when present); skipped when neither is.
"""
import os
import runpy
import sys
import types

import pytest
import torch

import fly_bproject_b200 as fb
from oracle import fly_oracle as fo
from oracle import ref_harness
from tests.helpers import ATOL_ENV, RTOL_ENV, OraclePool, assert_obs_close, oracle_tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_reference_ppo_drives_the_dropin_env():
    if not ref_harness.available():
        pytest.skip("reference sources not staged (oracle/_ref)")
    n = 20480                                   # ppo.py:120-122: mini_chunk_size 2, rollout_size 32
    ppo = ref_harness.load_ppo(fb.Fly)          # the reference module with its `Fly` name bound to the drop-in
    args = fb.make_args(num_envs=n, sim_device=DEV, state_pool=2, seed=21)
    torch.manual_seed(0)
    agent = ppo.PPO(args)
    assert type(agent).__module__ == "ppo" and isinstance(agent.env, fb.Fly)
    assert agent.rollout_size == 32 and agent.all_obs.shape == (32, n, 73) and agent.all_obs.is_cuda
    env = agent.env
    t = oracle_tables("limited_18", n, True)
    s = fo.EnvState(t)
    pool = OraclePool(env.state)
    pool.bind(s)
    for i in range(3):
        obs_before = s.obs_buf.clone()
        end = agent.run()                       # reference ppo.py:204-264, calling Fly.step of the drop-in
        torch.cuda.synchronize()
        assert end is False
        act = agent.all_acts[i].cpu()           # what the reference's policy sampled (clipped), on its own RNG
        assert float(act.abs().max()) <= 1.0 and act.shape == (n, 18)
        torch.testing.assert_close(agent.all_obs[i].cpu(), obs_before, rtol=RTOL_ENV, atol=1e-5)
        scaled, targets = fo.scale_actions(t, act)
        s.actions = scaled
        fo.apply_reset(t, s)
        pool.simulate(s)
        fo.post_physics(t, s)
        assert torch.equal(env.reset_buf.cpu(), s.reset_buf)
        assert torch.equal(env.progress_buf.cpu(), s.progress_buf)
        assert_obs_close(agent.all_next_obs[i].cpu(), s.obs_buf, 18)
        torch.testing.assert_close(agent.all_reward[i].squeeze(-1).cpu(), s.reward_buf, rtol=RTOL_ENV, atol=ATOL_ENV)
        torch.testing.assert_close(env.dof_targets.cpu(), targets, rtol=1e-6, atol=1e-6)
        assert torch.equal(agent.all_done.squeeze(-1).cpu(), 1 - s.reset_buf)          # ppo.py:230
    assert agent.mini_batch_number == 3 and agent.run_step == 3


def test_reference_trainer_drives_the_dropin_agent(monkeypatch, capsys):
    if not ref_harness.available():
        pytest.skip("reference sources not staged (oracle/_ref)")
    trainer = os.path.join(ref_harness.REFERENCE_ROOT, "trainer.py")
    made = []

    class BoundedPPO(fb.PPO):
        """trainer.py loops until the viewer is closed (`env.end`); headless, the test closes it after one rollout."""

        def __init__(self, args):
            super().__init__(args)
            made.append(self)

        def run(self):
            end = super().run()
            return end or self.run_step >= self.rollout_size

    fake = types.ModuleType("ppo")
    fake.PPO = BoundedPPO
    monkeypatch.setitem(sys.modules, "ppo", fake)          # `from ppo import PPO` (trainer.py:1) -> the drop-in
    monkeypatch.setattr(sys, "argv", ["trainer.py", "--num_envs", "2560", "--sim_device", DEV, "--headless", "True"])
    runpy.run_path(trainer, run_name="__main__")           # parses its args, builds the policy, main(), save, video, exit
    assert len(made) == 1
    agent = made[0]
    assert agent.num_envs == 2560 and agent.rollout_size == 256 and agent.mini_chunk_size == 16      # ppo.py:120-122
    assert agent.run_step == 256 and agent.optim_step == 75 and agent.mini_batch_number == 0
    assert bool(torch.isfinite(agent.net.flat).all())
    assert float(agent.action_var[0]) == pytest.approx(0.2 - 256 * 0.00001, rel=1e-5)                # ppo.py:236-237
    out = capsys.readouterr().out
    assert "Training" in out and "Steps: 0000 | Opt Step: 0000" in out                              # ppo.py:243, 258
