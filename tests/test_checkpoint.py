"""Checkpoint format: the reference's `.pth` layout (ppo.py:147-149, 266-273) + the resume sidecar."""
import os
from collections import OrderedDict

import pytest
import torch

import fly_bproject_b200 as fb
from fly_bproject_b200 import checkpoint as ck
from oracle import fly_oracle as fo
from tests.helpers import load_golden

DEV = "cuda:0"


def _reference_layout_file(tmp_path, g):
    """A file exactly as the reference writes it: torch.save(net.state_dict()) of an nn.Sequential-named module."""
    sd = OrderedDict((k, g["w0/" + k].clone()) for k in fo.PARAM_ORDER)
    path = str(tmp_path / "save1_standing_18dofs_7800.pth")
    torch.save(sd, path)
    return path, sd


def test_reference_layout_file_loads_and_validates(tmp_path):
    g = load_golden("ppo_n64_t32.npz")
    path, sd = _reference_layout_file(tmp_path, g)
    got, side = ck.load_checkpoint(path, 73, 18)
    assert side is None and list(got) == ck.PARAM_NAMES == fo.PARAM_ORDER
    for k in ck.PARAM_NAMES:
        assert torch.equal(got[k], sd[k])
    # the same file loads into the reference's own nn.Module layout (what `net.load_state_dict` wants)
    mods = torch.nn.ModuleDict({
        "shared_net": torch.nn.Sequential(torch.nn.Linear(73, 256), torch.nn.ELU(), torch.nn.Linear(256, 128), torch.nn.ELU()),
        "to_mean": torch.nn.Sequential(torch.nn.Linear(128, 64), torch.nn.ELU(), torch.nn.Linear(64, 18), torch.nn.ELU()),
        "to_value": torch.nn.Sequential(torch.nn.Linear(128, 64), torch.nn.ELU(), torch.nn.Linear(64, 1))})
    assert str(mods.load_state_dict(torch.load(path))) == "<All keys matched successfully>"
    with pytest.raises(RuntimeError, match="size mismatch"):
        ck.load_checkpoint(path, 145, 42)
    bad = OrderedDict(sd)
    bad.pop("to_value.2.bias")
    torch.save(bad, str(tmp_path / "bad.pth"))
    with pytest.raises(RuntimeError, match="missing"):
        ck.load_checkpoint(str(tmp_path / "bad.pth"))


def test_sidecar_round_trip_and_staleness(tmp_path):
    g = load_golden("ppo_n64_t32.npz")
    sd = OrderedDict((k, g["w3/" + k].clone()) for k in fo.PARAM_ORDER)
    path = str(tmp_path / "ck.pth")
    total = 70000
    extras = {"exp_avg": torch.randn(total), "exp_avg_sq": torch.rand(total), "adam_step": 3,
              "action_var": torch.full((18,), 0.19997), "optim_step": 3, "run_step": 96, "rng_counter": 96}
    ck.save_checkpoint(path, sd, extras)
    assert os.path.exists(ck.sidecar_path(path))
    # the .pth itself stays a plain reference file: 12 tensors, nothing else
    raw = torch.load(path)
    assert list(raw) == fo.PARAM_ORDER and all(isinstance(v, torch.Tensor) for v in raw.values())
    got, side = ck.load_checkpoint(path)
    for k in ("exp_avg", "exp_avg_sq", "action_var"):
        assert torch.equal(side[k], extras[k])
    assert (side["adam_step"], side["optim_step"], side["run_step"], side["rng_counter"]) == (3, 3, 96, 96)
    assert ck.load_checkpoint(path, want_sidecar=False)[1] is None
    ck.save_checkpoint(path, sd, None)                     # weights rewritten without extras: the stale sidecar goes
    assert not os.path.exists(ck.sidecar_path(path))
    with pytest.raises(ValueError):
        ck.save_checkpoint(path, sd, {"exp_avg": torch.zeros(1)})


def test_real_reference_checkpoint_loads_if_staged():
    from oracle import ref_harness
    path = os.path.join(ref_harness.REFERENCE_ROOT, "saves/save7_1_23/save1_standing_18dofs_7800.pth")
    if not os.path.isfile(path):
        pytest.skip("reference checkpoint not present")
    sd, side = ck.load_checkpoint(path, 73, 18)
    g = load_golden("ppo_n64_t32.npz")
    assert side is None
    for k in fo.PARAM_ORDER:
        assert torch.equal(sd[k], g["w0/" + k])            # the golden w0 IS this checkpoint


@pytest.mark.gpu
def test_args_load_reads_reference_layout_file(tmp_path):
    g = load_golden("ppo_n64_t32.npz")
    path, _ = _reference_layout_file(tmp_path, g)
    agent = fb.PPO(fb.make_args(num_envs=256, sim_device=DEV, horizon=16, verbose=False, load=True, load_path=path))
    torch.testing.assert_close(agent.net.pi(g["net_x"].to(DEV)).cpu(), g["net_pi"], rtol=1e-5, atol=1e-5)
    assert agent.optim_step == 0 and float(agent.action_var[0]) == pytest.approx(0.2)     # no sidecar: fresh, as the reference


@pytest.mark.gpu
def test_resume_is_bit_identical_to_uninterrupted_run(tmp_path):
    """1 update -> save -> fresh PPO with args.load -> 2nd update == uninterrupted 2nd update, bit for bit (parameters,
    Adam moments, variance, counters).  Both runs collect the second rollout from a fresh env of the same seed."""
    base = str(tmp_path / "run_")
    mk = lambda **kw: fb.make_args(num_envs=1024, sim_device=DEV, horizon=16, verbose=False, state_pool=2, seed=3,
                                   save=True, save_path=base, **kw)
    torch.manual_seed(0)
    a = fb.PPO(mk())
    a.collect_rollout(use_graph=False)
    a.update()
    a.save("75")
    a.env = fb.Fly(mk())
    a.collect_rollout(use_graph=False)
    a.update()
    torch.manual_seed(12345)                                # different init: everything must come from the files
    b = fb.PPO(mk(load=True, load_path=base + "75.pth"))
    assert b.optim_step == 75 and b.run_step == 16 and int(b.adam_step.item()) == 75
    assert int(b._rng_counter.item()) == 16
    b.collect_rollout(use_graph=False)
    b.update()
    torch.cuda.synchronize()
    assert torch.equal(a.all_acts, b.all_acts) and torch.equal(a.all_log_prob, b.all_log_prob)
    for name in ("exp_avg", "exp_avg_sq", "action_var", "adam_step"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    assert torch.equal(a.net.flat, b.net.flat)
    assert a.optim_step == b.optim_step == 150 and a.run_step == b.run_step == 32
    # without the sidecar the same file is a plain reference checkpoint: weights only
    c = fb.PPO(mk(load=True, load_path=base + "75.pth", resume=False))
    assert c.optim_step == 0 and float(c.action_var[0]) == pytest.approx(0.2) and int(c.adam_step.item()) == 0
    # evaluation (trainer.py --testing --load_path, ppo.py:151): weights from the file, variance 0.01 and fresh counters
    # even though a sidecar with the training variance sits next to it
    assert os.path.exists(ck.sidecar_path(base + "75.pth"))
    d = fb.PPO(mk(load=True, load_path=base + "75.pth", testing=True))
    assert torch.equal(d.action_var, torch.full_like(d.action_var, 0.01))
    assert d.optim_step == 0 and d.run_step == 0 and int(d.adam_step.item()) == 0
    e = fb.PPO(mk(load=True, load_path=base + "75.pth", resume=False))
    assert torch.equal(d.net.flat, e.net.flat)
    d.run()
    assert torch.equal(d.action_var, torch.full_like(d.action_var, 0.01))      # testing mode does not anneal
