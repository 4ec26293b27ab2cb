"""Constants and index tables of the reference env, packed for the C ABI.

Numbers come from the reference: literals of ``Fly.__init__`` (reference fly.py:16-51), joint limits / link order /
default pose extracted from its URDF + YAML assets by ``tools/make_tables.py`` into ``data/nmf_tables.json``
(Isaac Gym's own DOF/body ordering is not knowable offline — SURVEY.md §8a — so the tables are runtime data, in URDF
file order for synthetic runs; a live Isaac Gym binding would fill them from ``gym.find_actor_dof_index``).
"""
from __future__ import annotations

import json
import os
from typing import Optional, Sequence

from ._lib import EnvCfg, MAX_DOF

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "nmf_tables.json")

# reference fly.py:16-51
DT = 1.0 / 60.0
STARTING_HEIGHT = 2.0
MAX_EPISODE_LENGTH = 1500
DOF_VEL_SCALE = 0.2
HEADING_WEIGHT = 0.5
UP_WEIGHT = 0.75
ACTIONS_COST_SCALE = 0.005
ENERGY_COST_SCALE = 0.005
ENERGY_COST_SCALE_LOWGRAV = 1.0        # reference flyLowGrav.py:44
JOINTS_AT_LIMIT_COST_SCALE = 0.1
DEATH_COST = -2.0
TERMINATION_HEIGHT = 1.1
TERMINATION_HEIGHT_UP = 6.0
TARGET = (1000.0, 0.0, 0.0)


def load_asset(asset: str = "limited_18") -> dict:
    with open(_DATA) as f:
        return json.load(f)["assets"][asset]


def make_env_cfg(num_envs: int, asset: str = "limited_18", full_actions: Optional[bool] = None,
                 walking: bool = False, action_index: Optional[Sequence[int]] = None,
                 variant: str = "standard") -> EnvCfg:
    """Build the FlybEnvCfg for `num_envs` envs of the named asset.

    full_actions: actuate every DOF of the asset (A = D) instead of the reference's 18 named joints
    (reference fly.py:23-31).  Defaults to True only when the asset has exactly those 18 DOFs.
    variant: "standard" (fly.py) or "lowgrav" (flyLowGrav.py: energy_cost_scale = 1, line 44; its gravity, substeps,
    friction and joint stiffness, lines 36-37, 148-151, 222, configure PhysX, which is not part of this path).
    """
    if variant not in ("standard", "lowgrav"):
        raise ValueError(f"variant {variant!r}: expected 'standard' or 'lowgrav'")
    t = load_asset(asset)
    d = t["num_dof"]
    if d > MAX_DOF:
        raise ValueError(f"asset has {d} DOFs > FLYB_MAX_DOF={MAX_DOF}")
    if action_index is None:
        if full_actions is None:
            full_actions = d == 18
        action_index = list(range(d)) if full_actions else t["action_index_18"]
    aidx = sorted(int(i) for i in action_index)                      # reference fly.py:288-290 sorts them
    cfg = EnvCfg()
    cfg.num_envs, cfg.num_dof, cfg.num_act, cfg.num_bodies = num_envs, d, len(aidx), t["num_bodies"]
    for k, i in enumerate(aidx):
        cfg.action_idx[k] = i
    for k, i in enumerate(sorted(t["abdomen_local"])):
        cfg.abdomen_local[k] = i
    for k, i in enumerate(sorted(t["leg_tip_local"])):
        cfg.legtip_local[k] = i
    for i in range(d):
        cfg.lower[i] = t["lower"][i]
        cfg.upper[i] = t["upper"][i]
        cfg.initial_pos[i] = t["initial_pos"][i]
        cfg.initial_vel[i] = 0.0
    origin = [0.0] * 13                                              # reference fly.py:338-348
    origin[2] = STARTING_HEIGHT
    origin[6] = 1.0
    for i, v in enumerate(origin):
        cfg.origin_root[i] = v
    for i, v in enumerate(TARGET):
        cfg.target[i] = v
    for i, v in enumerate((-0.0, -0.0, -0.0, 1.0)):                  # quat_conjugate([0,0,0,1]) (fly.py:129)
        cfg.inv_start_rot[i] = v
    for i, v in enumerate((1.0, 0.0, 0.0)):
        cfg.basis0[i] = v
    for i, v in enumerate((0.0, 0.0, 1.0)):
        cfg.basis1[i] = v
    cfg.dt = DT
    cfg.dof_vel_scale = DOF_VEL_SCALE
    cfg.heading_weight = HEADING_WEIGHT
    cfg.up_weight = UP_WEIGHT
    cfg.actions_cost_scale = ACTIONS_COST_SCALE
    cfg.energy_cost_scale = ENERGY_COST_SCALE_LOWGRAV if variant == "lowgrav" else ENERGY_COST_SCALE
    cfg.joints_at_limit_cost_scale = JOINTS_AT_LIMIT_COST_SCALE
    cfg.death_cost = DEATH_COST
    cfg.termination_height = TERMINATION_HEIGHT
    cfg.termination_height_up = TERMINATION_HEIGHT_UP
    cfg.max_episode_length = float(MAX_EPISODE_LENGTH)
    cfg.walking = 1 if walking else 0
    return cfg
