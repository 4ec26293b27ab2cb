"""GPU: `flyb_env_post_physics_static` — the post-physics launch for hosts that cannot allocate device memory.

Some users of libflyb200.so have nothing to allocate with: a plain-ctypes / C binding that only forwards the simulator's
wrapped tensors (no torch, no CUDA-runtime binding), or code that runs inside the simulator's step loop, where cudaMalloc
would synchronise the device or fail because the simulator reserved the memory.  For them the kernel's cross-CTA scratch
(a counter + one reset mask per 64 envs) sits in an array compiled into the library, good for up to
FLYB_STATIC_MAX_ENVS envs and shared by every call on a device.

What is checked, through the C ABI: on the same inputs the entry point returns exactly what `flyb_env_post_physics` (with a
caller-owned scratch buffer) returns — masks, counters, compacted ids, obs, reward and potentials bit for bit (it is the
same kernel), `reward_sum` at the suite's bar for that atomically accumulated scalar — on TWO calls in a row that differ
in env count and reset pattern, so that whatever the first call left in the shared array cannot leak into the second;
then once more at the first size.  Env counts above the limit are refused with a message that names it.

Collected last on purpose (file name): `pytest -x` stops at the first failure, and a problem in this optional entry
point must not hide the parity tests.  It builds all its state itself, so it passes alone in a fresh process as well.
"""
import ctypes as C

import pytest
import torch

import fly_bproject_b200 as fb
from fly_bproject_b200 import _lib
from oracle import fly_oracle as fo
from tests.helpers import oracle_tables

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
STATIC_MAX_ENVS = 262144          # FLYB_STATIC_MAX_ENVS in include/flyb200.h

OUT = ["potentials", "prev_potentials", "progress", "reset", "obs", "reward", "up_vec", "heading_vec", "reset_ids",
       "reset_count", "obs_copy", "reward_copy", "reward_sum"]


def _buffers(n, asset, seed, flagged):
    """Inputs of one post-physics call (seeded synthetic physics state) + zeroed outputs, on the GPU."""
    t = oracle_tables(asset, n, True)
    st = fo.synth_state(t, seed=seed)
    g = torch.Generator().manual_seed(seed)
    cfg = fb.make_env_cfg(n, asset=asset, full_actions=True)
    num_obs = 19 + 3 * cfg.num_act
    f32 = dict(dtype=torch.float32, device=DEV)
    scaled, _ = fo.scale_actions(t, st["actions"])
    b = dict(
        root=st["root"].to(DEV).contiguous(), dof_state=st["dof_states"].to(DEV).contiguous(),
        force=st["force"].to(DEV).contiguous(), actions_scaled=scaled.to(DEV).contiguous(),
        potentials=(-torch.rand(n, generator=g) * 60000.0).to(DEV), prev_potentials=torch.zeros(n, **f32),
        progress=st["progress"].to(DEV).contiguous(),
        reset=(torch.rand(n, generator=g) < flagged).long().to(DEV),
        obs=torch.zeros(n, num_obs, **f32), reward=torch.zeros(n, **f32), up_vec=torch.zeros(n, 3, **f32),
        heading_vec=torch.zeros(n, 3, **f32), reset_ids=torch.full((n,), -1, dtype=torch.int32, device=DEV),
        reset_count=torch.full((1,), -1, dtype=torch.int32, device=DEV), obs_copy=torch.zeros(n, num_obs, **f32),
        reward_copy=torch.zeros(n, **f32), reward_sum=torch.zeros(1, **f32))
    return cfg, b


def _call(cfg, b, scratch):
    """scratch = a caller-owned buffer -> flyb_env_post_physics; None -> the entry point that needs none."""
    p, lib = _lib.ptr, _lib.lib()
    head = [p(b[k]) for k in ("root", "dof_state", "force", "actions_scaled", "potentials", "prev_potentials",
                               "progress", "reset", "obs", "reward", "up_vec", "heading_vec", "reset_ids",
                               "reset_count")]
    tail = [p(b["obs_copy"]), p(b["reward_copy"]), p(b["reward_sum"]), None, _lib.current_stream()]
    if scratch is None:
        _lib.check(lib.flyb_env_post_physics_static(C.byref(cfg), *head, *tail), "flyb_env_post_physics_static")
    else:
        _lib.check(lib.flyb_env_post_physics(C.byref(cfg), *head, p(scratch), *tail), "flyb_env_post_physics")


def _same(got, want, n):
    torch.cuda.synchronize()
    for k in OUT:
        if k == "reward_sum":       # one atomicAdd per warp: the order differs from run to run (bar of test_gpu_env.py)
            torch.testing.assert_close(got[k], want[k], rtol=1e-4, atol=1e-3)
        elif k == "reset_ids":
            cnt = int(want["reset_count"].item())
            assert torch.equal(got[k][:cnt], want[k][:cnt]), k
        else:
            assert torch.equal(got[k], want[k]), k
    # and the ids are what the mask says (== reset_buf.nonzero(), reference fly.py:447)
    ids = got["reset"].nonzero().squeeze(-1).to(torch.int32)
    assert int(got["reset_count"].item()) == ids.numel()
    assert torch.equal(got["reset_ids"][:ids.numel()], ids)


def test_static_scratch_entry_point_matches_the_allocating_one():
    # second call: fewer envs, a ragged last CTA and a different mask -> stale words of call 1 sit behind its bitmap
    for n, asset, seed, flagged in [(16384, "limited_18", 11, 0.3), (1001, "full_42", 12, 0.9),
                                    (16384, "limited_18", 13, 0.0)]:
        cfg, b = _buffers(n, asset, seed, flagged)
        want = {k: v.clone() for k, v in b.items()}
        own = torch.zeros(_lib.lib().flyb_env_scratch_bytes(n), dtype=torch.uint8, device=DEV)
        _call(cfg, want, own)
        _call(cfg, b, None)
        _same(b, want, n)
        assert int(b["reset_count"].item()) > 0


def test_static_scratch_entry_point_refuses_too_many_envs():
    cfg = fb.make_env_cfg(STATIC_MAX_ENVS + 1, asset="limited_18", full_actions=True)   # refused before any launch
    rc = _lib.lib().flyb_env_post_physics_static(C.byref(cfg), *([None] * 17), None, None)
    assert rc == -1
    msg = _lib.lib().flyb_last_error().decode()
    assert "FLYB_STATIC_MAX_ENVS" in msg and str(STATIC_MAX_ENVS) in msg
