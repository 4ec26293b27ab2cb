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


_walking = {}


def load_walking_reward(leg_term: bool = True):
    """``compute_fly_reward2`` with the reference's OWN commented-out walking total switched back on.

    reference fly.py:747-748 keeps the walking total as two ``#`` lines; fly.py:750 is the live standing total.  The
    function's source is taken from the reference file as it lies, the ``#`` in front of those two lines is removed, the
    standing total line is dropped, and the result is compiled by ``torch.jit.script`` — every other line of the function
    runs as written.  Returns the scripted function.

    Two readings, both pinned (fixtures env_walking_d18_a18_n211.npz / env_walking_literal_d18_a18_n211.npz):
    ``leg_term=False`` is the LITERAL un-commenting: inside the dead line the author had commented the leg term out once
    more (``#+ leg_ground_reward``), so as written the walking total has no leg term.
    ``leg_term=True`` (what ``walking=True`` means in the oracle and in the CUDA kernels) also re-enables that nested
    ``#+ leg_ground_reward``.  This is an INTERPRETATION, not code the reference can execute as written: it follows
    BASELINE.json configs[2] / the project report (p.7), which list the leg-contact term in the walking reward, and the
    live standing total (fly.py:750), which does add it.
    """
    global _walking
    if leg_term not in _walking:
        load_fly()                                          # stubs in place, reference importable
        with open(os.path.join(REFERENCE_ROOT, "fly.py")) as f:
            lines = f.read().split("\n")
        start = next(i for i, ln in enumerate(lines) if ln.startswith("def compute_fly_reward2("))
        end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith("def ") or lines[i].startswith("@"))
        body, out, seen = lines[start:end], [], 0
        for ln in body:
            st = ln.strip()
            if st.startswith("#total_reward = progress_reward * 2"):
                out.append("    " + st[1:]); seen += 1
            elif st.startswith("#") and "actions_cost_scale * actions_cost" in st and seen == 1:
                cont = st[1:].strip()
                if "#+ leg_ground_reward" not in cont:
                    raise RuntimeError("reference fly.py:748 does not end in the nested '#+ leg_ground_reward'")
                cont = cont.replace("#+ leg_ground_reward", "+ leg_ground_reward" if leg_term else "")
                out.append("        " + cont.rstrip()); seen += 1
            elif st.startswith("total_reward = alive_reward + up_reward * orient_reward") and seen == 2:
                seen += 1                                   # the standing total (fly.py:750) gives way
            else:
                out.append(ln)
        if seen != 3:
            raise RuntimeError("reference fly.py:747-750 does not look the way this edit expects")
        src = "\n".join(out).replace("def compute_fly_reward2(", "def compute_fly_reward_walking(")
        # torch.jit.script needs the source on disk (inspect.getsource): a temp module file
        import tempfile
        import importlib.util
        d = tempfile.mkdtemp(prefix="flyb_walking_")
        path = os.path.join(d, "fly_walking_total.py")
        with open(path, "w") as f:
            f.write("import torch\nfrom typing import Tuple\n\n\n" + src + "\n")
        spec = importlib.util.spec_from_file_location("fly_walking_total", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _walking[leg_term] = torch.jit.script(mod.compute_fly_reward_walking)
    return _walking[leg_term]


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
