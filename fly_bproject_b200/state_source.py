"""Where the three Isaac Gym state tensors come from.

The reference obtains zero-copy CUDA views of PhysX state with ``gym.acquire_*_tensor`` + ``gymtorch.wrap_tensor``
(reference fly.py:378-395) and refreshes them after ``gym.simulate`` (fly.py:401-403, 482-485).  The PhysX articulation
step is not reimplemented here (BASELINE.json north_star) and Isaac Gym is not installable offline, so the default
source synthesises those tensors at the named shapes: ``root (N,13)``, ``dof_state (N,D,2)``, ``net_contact_force
(N*B,3)``, with the value distributions of SURVEY.md §8(d) that exercise every termination branch.

`SyntheticStateSource` holds a POOL of pre-generated states and `simulate()` rotates through it (a pointer swap, no
copy), so that consecutive steps read different HBM lines; with pool bytes > L2 (126 MB) every step streams its
state from HBM.  Writes the env makes into the live state (reset rows, fly.py:454-459) land in the pool entry; a real
physics step would overwrite them, so every time the rotation wraps around the root / DOF slabs of all entries are
restored from a pristine copy (two flat device copies per `pool` steps) — otherwise every row that ever triggered a
reset would stay healed and the termination mix would drift from ~50 % of the envs to the abdomen hits and time-outs.
"""
from __future__ import annotations

import math
from typing import List

import torch


def synth_state(cfg, seed: int, device, progress: bool = False):
    """One seeded synthetic state (same generator as oracle/fly_oracle.py:synth_state so fixtures agree)."""
    g = torch.Generator().manual_seed(seed)
    n, d, b = cfg.num_envs, cfg.num_dof, cfg.num_bodies
    lower = torch.tensor(list(cfg.lower)[:d])
    upper = torch.tensor(list(cfg.upper)[:d])

    def U(*shape):
        return torch.rand(*shape, generator=g)

    def Nrm(*shape):
        return torch.randn(*shape, generator=g)

    root = torch.zeros(n, 13)
    root[:, 0:2] = U(n, 2) * 10 - 5
    sel = U(n)
    z = 2.0 + 0.3 * Nrm(n)
    z = torch.where(sel < 0.10, 0.5 + 0.6 * U(n), z)
    z = torch.where((sel >= 0.10) & (sel < 0.15), 1.1 + 0.3 * U(n), z)
    z = torch.where((sel >= 0.15) & (sel < 0.20), 6.0 + U(n), z)
    root[:, 2] = z
    kind = U(n)
    yaw = U(n) * 2 * math.pi - math.pi
    tilt = 0.05 * Nrm(n, 2)
    q_yaw = torch.stack([tilt[:, 0], tilt[:, 1], torch.sin(yaw / 2), torch.cos(yaw / 2)], dim=-1)
    q_rand = Nrm(n, 4)
    q_flip = torch.stack([torch.cos(yaw / 2), torch.sin(yaw / 2), 0.3 * Nrm(n), 0.3 * Nrm(n)], dim=-1)
    q = torch.where((kind < 0.7).unsqueeze(-1), q_yaw, torch.where((kind < 0.9).unsqueeze(-1), q_rand, q_flip))
    root[:, 3:7] = q / q.norm(dim=-1, keepdim=True)
    root[:, 7:13] = Nrm(n, 6)

    dof = torch.zeros(n, d, 2)
    dof[..., 0] = lower + (upper - lower) * U(n, d)
    dof[..., 1] = 2.0 * Nrm(n, d)

    force = Nrm(n * b, 3)
    force = torch.where((U(n * b) < 0.7).unsqueeze(-1), torch.zeros_like(force), force)
    ab_local = torch.tensor(sorted(cfg.abdomen_local))
    ab_ids = (torch.arange(n).view(n, 1) * b + ab_local.view(1, -1)).reshape(-1)
    ab = force[ab_ids].view(n, -1, 3)
    ab = torch.where((U(n) < 0.85).view(n, 1, 1), torch.zeros_like(ab), ab)
    force[ab_ids] = ab.view(-1, 3)
    out = {"root": root.to(device), "dof_state": dof.view(n, 2 * d).contiguous().to(device),
           "force": force.to(device)}
    if progress:
        out["progress"] = torch.randint(1, 1601, (n,), generator=g).to(device)
    return out


class SyntheticStateSource:
    """Pool of synthetic Isaac Gym states; `simulate()` advances to the next pool entry."""

    def __init__(self, cfg, device, pool: int = 4, seed: int = 1234, restore: bool = True):
        self.cfg = cfg
        self.device = torch.device(device)
        npool = max(1, pool)
        states = [synth_state(cfg, seed + 7919 * i, self.device) for i in range(npool)]
        # root / dof_state of all entries live in ONE slab each (entries are views), so that the restore is two copies;
        # every entry starts on a 256-byte boundary whatever N is (the kernels take 16-byte aligned state tensors)
        def slab(key):
            numel = states[0][key].numel()
            pitch = (numel + 63) // 64 * 64
            flat = torch.zeros(npool * pitch, dtype=torch.float32, device=self.device)
            views = [flat[i * pitch:i * pitch + numel].view(states[0][key].shape) for i in range(npool)]
            for v, s in zip(views, states):
                v.copy_(s[key])
            return flat, views
        self._root_all, roots = slab("root")
        self._dof_all, dofs = slab("dof_state")
        self.pool: List[dict] = [{"root": roots[i], "dof_state": dofs[i], "force": states[i]["force"]}
                                 for i in range(npool)]
        self.restore = bool(restore)
        self._root_pristine = self._root_all.clone() if self.restore else None
        self._dof_pristine = self._dof_all.clone() if self.restore else None
        self.index = 0

    @property
    def pool_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for s in self.pool for t in s.values())

    @property
    def root(self) -> torch.Tensor:
        return self.pool[self.index]["root"]

    @property
    def dof_state(self) -> torch.Tensor:
        return self.pool[self.index]["dof_state"]

    @property
    def force(self) -> torch.Tensor:
        return self.pool[self.index]["force"]

    def set_dof_position_targets(self, targets: torch.Tensor) -> None:
        """gym.set_dof_position_target_tensor (fly.py:657): consumed by PhysX; nothing to do synthetically."""

    def set_reset_ids(self, ids_int32: torch.Tensor, count_dev: torch.Tensor) -> None:
        """gym.set_*_tensor_indexed (fly.py:462-468) take the compacted int32 ids; count stays on the device."""

    def simulate(self) -> None:
        """gym.simulate + fetch_results + refresh_* (fly.py:482-485, 401-403)."""
        self.index = (self.index + 1) % len(self.pool)
        if self.index == 0 and self.restore:      # the "physics" overwrites what Fly.reset wrote into the state
            self._root_all.copy_(self._root_pristine)
            self._dof_all.copy_(self._dof_pristine)


class TensorStateSource:
    """Caller-owned state tensors (e.g. gymtorch.wrap_tensor views, or fixtures in tests); simulate() is a hook."""

    def __init__(self, root, dof_state, force, simulate_fn=None):
        self.root, self.dof_state, self.force = root, dof_state, force
        self._simulate = simulate_fn

    def set_dof_position_targets(self, targets):
        pass

    def set_reset_ids(self, ids_int32, count_dev):
        pass

    def simulate(self):
        if self._simulate is not None:
            self._simulate(self)
