"""Shared helpers of the test-suite (oracle-side table construction, golden loading, tolerances)."""
import json
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
TABLES = json.load(open(os.path.join(ROOT, "fly_bproject_b200", "data", "nmf_tables.json")))["assets"]

# north_star tolerances: obs / reward / GAE within 1e-5 relative (fp32); one PPO step within 1e-4 relative.
RTOL_ENV = 1e-5
ATOL_ENV = 2e-6     # absolute floor for values near zero (fp32 eps at O(1..10) magnitudes)
RTOL_PPO = 1e-4
TWO_PI = 2 * np.pi
SEAM = 1e-3         # angles within this distance of the 0 / 2*pi seam are compared modulo 2*pi


def oracle_tables(asset: str, n: int, full_actions: bool):
    from oracle import fly_oracle as fo
    t = TABLES[asset]
    aidx = list(range(t["num_dof"])) if full_actions else t["action_index_18"]
    return fo.EnvTables(n, t["lower"], t["upper"], t["initial_pos"], aidx, t["abdomen_local"], t["leg_tip_local"],
                        t["num_bodies"])


def load_golden(name: str):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: torch.from_numpy(z[k]) for k in z.files}


ENV_CASES = [
    ("env_d18_a18_n257.npz", "limited_18", 257, True),
    ("env_d42_a18_n131.npz", "full_42", 131, False),
    ("env_d42_a42_n131.npz", "full_42", 131, True),
]


def angle_cols(num_act: int):
    """obs columns holding angles wrapped to [0, 2*pi): yaw, roll, pitch; angle_to_target inherits yaw's seam."""
    return [7, 8, 9, 12 + 3 * num_act]


def assert_obs_close(got: torch.Tensor, want: torch.Tensor, num_act: int, rtol=RTOL_ENV, atol=ATOL_ENV):
    """Compare observation rows; wrapped angles are compared modulo 2*pi when they sit on the seam."""
    got, want = got.double(), want.double()
    diff = (got - want).abs()
    seam_hits = 0
    for c in angle_cols(num_act):
        d = diff[:, c]
        wrapped = torch.minimum(d, (d - TWO_PI).abs())
        near = (want[:, c].abs() % TWO_PI < SEAM) | ((TWO_PI - want[:, c].abs() % TWO_PI) < SEAM) | (d > 6.0)
        seam_hits += int((near & (d > 1.0)).sum())
        diff[:, c] = torch.where(near, wrapped, d)
    tol = atol + rtol * want.abs()
    bad = diff > tol
    assert not bad.any(), (f"{int(bad.sum())} obs entries out of tolerance; worst abs diff {float(diff.max()):.3e} "
                           f"at {np.unravel_index(int(diff.argmax()), diff.shape)}")
    return seam_hits


class OraclePool:
    """CPU mirror of `SyntheticStateSource`: the same pool entries, the same rotation and the same restore of the
    root / DOF slabs whenever the rotation wraps around (what a real physics step overwriting the reset rows does)."""

    def __init__(self, source):
        self.entries = [{k: v.cpu().clone() for k, v in e.items()} for e in source.pool]
        self.pristine = [{k: e[k].clone() for k in ("root", "dof_state")} for e in self.entries]
        self.restore = bool(getattr(source, "restore", False))
        self.index = 0

    def bind(self, s):
        e = self.entries[self.index]
        s.root, s.dof_states, s.force = e["root"], e["dof_state"], e["force"]

    def simulate(self, s):
        self.index = (self.index + 1) % len(self.entries)
        if self.index == 0 and self.restore:
            for e, p0 in zip(self.entries, self.pristine):
                e["root"].copy_(p0["root"])
                e["dof_state"].copy_(p0["dof_state"])
        self.bind(s)
