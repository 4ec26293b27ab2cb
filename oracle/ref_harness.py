"""Import the reference's own ``fly.py`` / ``ppo.py`` unmodified, on CPU torch.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Usable where the reference's files
exist: the read-only checkout ``/root/reference`` in the build container, or the
byte-for-byte staged copy ``oracle/_ref`` (``oracle/make_ref.py``; git-ignored,
travels with the gpurun snapshot) on the GPU box.  It is used to (1) validate the
travelling restatement in ``oracle/fly_oracle.py``
(``tests/test_oracle_vs_reference.py``), (2) generate the golden vectors under
``tests/golden/`` (``tests/golden/make_golden.py``) and (3) time the reference's
unmodified code as the CPU arm of ``bench.py``.

Recipe: put stub modules for ``isaacgym*`` /
``isaacgymenvs*`` into ``sys.modules`` whose ``*``-exports are the restated
helpers of ``oracle/isaac_math.py`` (+ ``np``, which reference fly.py:65,212
relies on the star import for), then ``import fly`` → the two
``@torch.jit.script`` functions compile; swap ``fly.Fly`` for a caller-supplied
env class and ``import ppo`` → the reference ``Net`` / ``PPO`` run unmodified.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import numpy as np
import torch

def _find_root() -> str:
    env = os.environ.get("FLYB200_REFERENCE_ROOT")
    if env:
        return env
    staged = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
    for cand in ("/root/reference", staged):
        if os.path.isfile(os.path.join(cand, "fly.py")):
            return cand
    return "/root/reference"


REFERENCE_ROOT = _find_root()

_STUB_NAMES = [
    "isaacgym", "isaacgym.gymapi", "isaacgym.gymtorch", "isaacgym.torch_utils",
    "isaacgymenvs", "isaacgymenvs.utils", "isaacgymenvs.utils.torch_jit_utils",
]


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "fly.py"))


def _install_stubs() -> None:
    from . import isaac_math

    for name in _STUB_NAMES:
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    for name in _STUB_NAMES:
        if "." in name:
            parent, child = name.rsplit(".", 1)
            setattr(sys.modules[parent], child, sys.modules[name])
    for modname in ("isaacgym.torch_utils", "isaacgymenvs.utils.torch_jit_utils"):
        mod = sys.modules[modname]
        exported = list(isaac_math.EXPORTED) + ["np", "torch"]
        for sym in isaac_math.EXPORTED:
            setattr(mod, sym, getattr(isaac_math, sym))
        mod.np = np
        mod.torch = torch
        mod.__all__ = exported


_fly = None
_ppo = None
_lowgrav = None


def load_fly():
    """The reference ``fly`` module (its ``Fly`` class cannot be constructed: it needs gymapi)."""
    global _fly
    if _fly is None:
        if not available():
            raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
        _install_stubs()
        sys.dont_write_bytecode = True  # the reference tree is read-only
        if REFERENCE_ROOT not in sys.path:
            sys.path.insert(0, REFERENCE_ROOT)
        _fly = importlib.import_module("fly")
    return _fly


def load_fly_lowgrav():
    """The reference ``flyLowGrav`` module (same two TorchScript functions, low-gravity constants and step order)."""
    global _lowgrav
    if _lowgrav is None:
        load_fly()
        _lowgrav = importlib.import_module("flyLowGrav")
    return _lowgrav


def load_ppo(env_cls):
    """The reference ``ppo`` module with ``fly.Fly`` replaced by ``env_cls``.

    ``env_cls(args)`` must offer the duck-typed surface ppo.py uses
    (ppo.py:110-112, 210-211, 223, 228-230): ``num_act, num_obs, obs_buf,
    reward_buf, reset_buf, end, step(actions)``.
    """
    global _ppo
    fly = load_fly()
    fly.Fly = env_cls
    if _ppo is None:
        _ppo = importlib.import_module("ppo")
    _ppo.Fly = env_cls
    return _ppo
