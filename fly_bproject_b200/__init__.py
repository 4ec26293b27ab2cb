"""flyb200 — B200-native (sm_100a) implementation of fly_bProject's per-step hot path.

Drop-in classes behind the reference's own interface: `Fly` (reference fly.py) and `PPO` / `Net` (reference
ppo.py).  All tensor math runs in libflyb200.so (hand-written CUDA, C ABI in include/flyb200.h); importing this
package does not need a GPU, constructing `Fly` / `PPO` does — there is no CPU fallback.
"""
from ._lib import FlybError, lib  # noqa: F401
from .tables import make_env_cfg  # noqa: F401
from .env import Fly, make_args  # noqa: F401
from .agent import Net, PPO  # noqa: F401
from .state_source import SyntheticStateSource, TensorStateSource  # noqa: F401

__all__ = ["Fly", "PPO", "Net", "make_args", "make_env_cfg", "SyntheticStateSource", "TensorStateSource",
           "FlybError", "lib"]
