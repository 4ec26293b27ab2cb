"""Checkpoint files: the reference's `.pth` layout plus an opt-in sidecar with what the reference loses on resume.

The reference saves ONLY ``net.state_dict()`` (``torch.save(self.net.state_dict(), path)``, reference ppo.py:266-273;
written from trainer.py:48 and every ``save_freq`` optimiser steps, ppo.py:249-252) and loads it with
``net.load_state_dict(torch.load(path))`` (ppo.py:147-149): 12 tensors under the ``nn.Sequential`` names of
ppo.py:18-38.  A resumed run therefore restarts Adam's moments, the annealed action variance (back to 0.2, ppo.py:152)
and both step counters.  `save_checkpoint` keeps that file byte-compatible — the 95 live-architecture ``.pth`` files of
the reference's ``saves/`` load here and files written here load in the reference — and puts the rest next to it in
``<path>.flyb.pt``: ``exp_avg``, ``exp_avg_sq`` (flat, FlybNetLayout order), ``adam_step``,
``action_var``, ``optim_step``, ``run_step``, the Philox draw counter and the layout they were saved under.

Host-side only (no CUDA): tensors are moved to the CPU before they are written and come back as CPU tensors.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

PARAM_NAMES = [
    "shared_net.0.weight", "shared_net.0.bias", "shared_net.2.weight", "shared_net.2.bias",
    "to_mean.0.weight", "to_mean.0.bias", "to_mean.2.weight", "to_mean.2.bias",
    "to_value.0.weight", "to_value.0.bias", "to_value.2.weight", "to_value.2.bias",
]
SIDECAR_SUFFIX = ".flyb.pt"
SIDECAR_VERSION = 1
SIDECAR_KEYS = ("exp_avg", "exp_avg_sq", "adam_step", "action_var", "optim_step", "run_step", "rng_counter")


def param_shapes(num_obs: int, num_act: int):
    o, a = num_obs, num_act
    return [(256, o), (256,), (128, 256), (128,), (64, 128), (64,), (a, 64), (a,), (64, 128), (64,), (1, 64), (1,)]


def sidecar_path(path: str) -> str:
    return path + SIDECAR_SUFFIX


def validate_state_dict(sd, num_obs: Optional[int] = None, num_act: Optional[int] = None) -> Tuple[int, int]:
    """Check the 12 reference keys / shapes; returns (num_obs, num_act) as the file implies them."""
    missing = [k for k in PARAM_NAMES if k not in sd]
    unexpected = [k for k in sd if k not in PARAM_NAMES]
    if missing or unexpected:
        raise RuntimeError(f"Error(s) in loading state_dict for Net: missing {missing}, unexpected {unexpected}")
    o = int(sd["shared_net.0.weight"].shape[1])
    a = int(sd["to_mean.2.weight"].shape[0])
    if (num_obs is not None and o != num_obs) or (num_act is not None and a != num_act):
        raise RuntimeError(f"size mismatch: checkpoint is for num_obs={o}, num_act={a}; "
                           f"the model has num_obs={num_obs}, num_act={num_act}")
    for k, shp in zip(PARAM_NAMES, param_shapes(o, a)):
        if tuple(sd[k].shape) != shp:
            raise RuntimeError(f"size mismatch for {k}: checkpoint {tuple(sd[k].shape)} vs model {shp}")
    return o, a


def save_checkpoint(path: str, state_dict: Dict[str, torch.Tensor], extras: Optional[dict] = None) -> None:
    """`path`: the reference-format ``.pth`` (12 CPU tensors, an OrderedDict like ``nn.Module.state_dict()``).
    `extras` (optional): the sidecar's content (SIDECAR_KEYS); written to ``path + '.flyb.pt'``."""
    o, a = validate_state_dict(state_dict)
    sd = OrderedDict((k, state_dict[k].detach().to("cpu", torch.float32).contiguous()) for k in PARAM_NAMES)
    torch.save(sd, path)                                                       # ppo.py:266-273
    if extras is not None:
        missing = [k for k in SIDECAR_KEYS if k not in extras]
        if missing:
            raise ValueError(f"sidecar is missing {missing}")
        side = {"version": SIDECAR_VERSION, "num_obs": o, "num_act": a}
        for k in SIDECAR_KEYS:
            v = extras[k]
            side[k] = v.detach().cpu().clone() if isinstance(v, torch.Tensor) else int(v)
        torch.save(side, sidecar_path(path))
    elif os.path.exists(sidecar_path(path)):
        os.remove(sidecar_path(path))        # a stale sidecar must not be paired with new weights


def load_checkpoint(path: str, num_obs: Optional[int] = None, num_act: Optional[int] = None,
                    want_sidecar: bool = True) -> Tuple["OrderedDict[str, torch.Tensor]", Optional[dict]]:
    """Returns (state_dict on the CPU, sidecar dict or None).  A file without sidecar is a plain reference checkpoint:
    the caller starts Adam / variance / counters afresh, exactly as the reference does (ppo.py:147-155)."""
    sd = torch.load(path, map_location="cpu")                                 # ppo.py:149
    o, a = validate_state_dict(sd, num_obs, num_act)
    side = None
    sp = sidecar_path(path)
    if want_sidecar and os.path.exists(sp):
        side = torch.load(sp, map_location="cpu")
        if side.get("version") != SIDECAR_VERSION:
            raise RuntimeError(f"{sp}: sidecar version {side.get('version')} (this build reads {SIDECAR_VERSION})")
        if side["num_obs"] != o or side["num_act"] != a:
            raise RuntimeError(f"{sp}: sidecar is for num_obs={side['num_obs']}, num_act={side['num_act']}, "
                               f"the weights for {o}, {a}")
        missing = [k for k in SIDECAR_KEYS if k not in side]
        if missing:
            raise RuntimeError(f"{sp}: sidecar is missing {missing}")
    return OrderedDict((k, sd[k]) for k in PARAM_NAMES), side
