"""CPU: the numbers bench.py builds its roofline objects from are SURVEY.md §8(d)'s, and its configurations are
BASELINE.json's."""
import json
import os

import bench
from tests.helpers import ROOT


def test_work_model_matches_survey_8d():
    w = bench.work_model(18, 18)
    assert w["O"] == 73
    assert w["bytes_env"] == 764                    # 420 read + 344 written per env-step, kernel (a)
    assert w["bytes_step"] == 912                   # fused (a) + (b)
    assert w["flops_act"] == 121600                 # mu-MLP 73 -> 256 -> 128 -> 64 -> 18
    assert w["flops_update"] == 414336              # 6 * 69 056 weight MACs, trunk once
    assert w["bytes_gae"] == 20
    w42 = bench.work_model(42, 42)
    assert w42["O"] == 145 and w42["bytes_env"] == 1340 and w42["flops_act"] == 161536


def test_configs_are_the_baseline_ones():
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert len(base["configs"]) == 5
    c = bench.CONFIGS
    assert (c["c0"]["n"], c["c0"]["horizon"]) == (1000, 640)            # trainer.py defaults (ppo.py:118-122)
    assert (c["c1"]["n"], c["c1"]["horizon"]) == (16384, 32)            # the metric's config
    assert (c["c2"]["n"], c["c2"]["horizon"], c["c2"]["walking"]) == (8192, 80, True)     # 65536 over 8 GPUs
    assert (c["c3"]["n"], c["c3"]["horizon"], c["c3"]["asset"], c["c3"]["full"]) == (32768, 64, "full_42", True)
    assert c["c3b"]["full"] is False
    for k, v in c.items():
        assert v["horizon"] % v["pool"] == 0, k                         # the rollout graph needs T % pool == 0
    assert bench.METRIC == "env_steps_per_sec" and bench.UNIT == "env-steps/s"
