"""ctypes binding of libflyb200.so (the C ABI declared in include/flyb200.h).

The product path has no CPU implementation: `lib()` raises when the shared object is missing, and every compute
call raises `FlybError` when the library reports a failure (no CUDA device, bad argument, launch error).
"""
from __future__ import annotations

import ctypes as C
import os

MAX_DOF = 64
NUM_ABDOMEN = 5
NUM_LEGTIPS = 6

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libflyb200.so")


class FlybError(RuntimeError):
    pass


class EnvCfg(C.Structure):
    _fields_ = [
        ("num_envs", C.c_int32), ("num_dof", C.c_int32), ("num_act", C.c_int32), ("num_bodies", C.c_int32),
        ("action_idx", C.c_int32 * MAX_DOF),
        ("abdomen_local", C.c_int32 * NUM_ABDOMEN), ("legtip_local", C.c_int32 * NUM_LEGTIPS),
        ("lower", C.c_float * MAX_DOF), ("upper", C.c_float * MAX_DOF),
        ("initial_pos", C.c_float * MAX_DOF), ("initial_vel", C.c_float * MAX_DOF),
        ("origin_root", C.c_float * 13), ("target", C.c_float * 3), ("inv_start_rot", C.c_float * 4),
        ("basis0", C.c_float * 3), ("basis1", C.c_float * 3),
        ("dt", C.c_float), ("dof_vel_scale", C.c_float), ("heading_weight", C.c_float), ("up_weight", C.c_float),
        ("actions_cost_scale", C.c_float), ("energy_cost_scale", C.c_float),
        ("joints_at_limit_cost_scale", C.c_float), ("death_cost", C.c_float), ("termination_height", C.c_float),
        ("termination_height_up", C.c_float), ("max_episode_length", C.c_float), ("walking", C.c_int32),
    ]


MAX_PEERS = 8


class PeerGroup(C.Structure):
    """FlybPeerGroup: the ranks' gradient-exchange buffers as mapped in this process."""
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("total_pad", C.c_int32), ("timeout_ms", C.c_int32),
                ("base", C.c_void_p * MAX_PEERS)]


class StepTail(C.Structure):
    _fields_ = [("action_var", C.c_void_p), ("num_act", C.c_int32), ("decay", C.c_float), ("floor", C.c_float),
                ("step_counter", C.c_void_p), ("steps", C.c_int32)]


class StepIO(C.Structure):
    """FlybStepIO: buffers of one fused env-step launch (flyb_env_step_fused)."""
    _fields_ = [
        ("root", C.c_void_p), ("dof_state", C.c_void_p), ("force", C.c_void_p),
        ("actions_scaled", C.c_void_p), ("potentials", C.c_void_p), ("prev_potentials", C.c_void_p),
        ("progress", C.c_void_p), ("reset", C.c_void_p),
        ("obs", C.c_void_p), ("obs_ring", C.c_void_p), ("reward", C.c_void_p), ("reward_ring", C.c_void_p),
        ("reward_sum", C.c_void_p), ("up_vec", C.c_void_p), ("heading_vec", C.c_void_p),
        ("reset_ids", C.c_void_p), ("reset_count", C.c_void_p), ("scratch", C.c_void_p),
        ("action_var", C.c_void_p), ("noise", C.c_void_p), ("seed", C.c_uint64), ("offset", C.c_uint64),
        ("offset_dev", C.c_void_p), ("action", C.c_void_p), ("logp", C.c_void_p), ("dof_targets", C.c_void_p),
        ("anneal_steps", C.c_int32), ("anneal_decay", C.c_float), ("anneal_floor", C.c_float),
    ]


class NetLayout(C.Structure):
    _fields_ = [("num_obs", C.c_int32), ("num_act", C.c_int32), ("k0_pad", C.c_int32), ("total", C.c_int32),
                ("off", C.c_int32 * 12)]


class PpoCfg(C.Structure):
    _fields_ = [("clip", C.c_double), ("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double),
                ("eps", C.c_double), ("max_grad_norm", C.c_double)]


_P = C.c_void_p
# name -> (restype, argtypes); must list every symbol include/flyb200.h declares (tests check this)
SIGNATURES = {
    "flyb_abi_version": (C.c_int, []),
    "flyb_last_error": (C.c_char_p, []),
    "flyb_device_count": (C.c_int, []),
    "flyb_net_layout": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(NetLayout)]),
    "flyb_debug_set_trace": (C.c_int, [_P]),
    "flyb_wpack_bytes": (C.c_size_t, [C.POINTER(NetLayout)]),
    "flyb_pack_weights": (C.c_int, [C.POINTER(NetLayout), _P, _P, _P]),
    "flyb_env_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "flyb_env_post_physics": (C.c_int, [C.POINTER(EnvCfg)] + [_P] * 18 + [C.POINTER(StepTail), _P]),
    "flyb_env_post_physics_static": (C.c_int, [C.POINTER(EnvCfg)] + [_P] * 17 + [C.POINTER(StepTail), _P]),
    "flyb_env_pre_physics": (C.c_int, [C.POINTER(EnvCfg)] + [_P] * 10),
    "flyb_env_reset": (C.c_int, [C.POINTER(EnvCfg)] + [_P] * 7),
    "flyb_env_step_fused_supported": (C.c_int, [C.POINTER(NetLayout), C.POINTER(EnvCfg)]),
    "flyb_env_step_fused": (C.c_int, [C.POINTER(NetLayout), _P, _P, C.POINTER(EnvCfg), C.POINTER(StepIO), _P]),
    "flyb_policy_act": (C.c_int, [C.POINTER(NetLayout), _P, _P, _P, C.c_int64, _P, _P, C.c_uint64, C.c_uint64, _P,
                                  _P, _P, _P, _P, C.POINTER(EnvCfg)] + [_P] * 9),
    "flyb_net_forward": (C.c_int, [C.POINTER(NetLayout), _P, _P, _P, C.c_int64, _P, _P, _P]),
    "flyb_gae": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_float, C.c_float, _P, _P, _P]),
    "flyb_ppo_workspace_bytes": (C.c_size_t, [C.POINTER(NetLayout), C.c_int64]),
    "flyb_ppo_grad": (C.c_int, [C.POINTER(NetLayout), C.POINTER(PpoCfg)] + [_P] * 8 + [C.c_int64] + [_P] * 3
                      + [C.POINTER(PeerGroup), _P]),
    "flyb_ppo_apply": (C.c_int, [C.POINTER(NetLayout), C.POINTER(PpoCfg)] + [_P] * 5
                       + [C.c_float, _P, _P, _P, C.POINTER(PeerGroup), _P]),
    "flyb_ppo_step": (C.c_int, [C.POINTER(NetLayout), C.POINTER(PpoCfg)] + [_P] * 8 + [C.c_int64] + [_P] * 8),
    "flyb_ppo_step_dp": (C.c_int, [C.POINTER(NetLayout), C.POINTER(PpoCfg)] + [_P] * 8 + [C.c_int64] + [_P] * 5
                         + [C.c_float, _P, _P, C.POINTER(PeerGroup), _P]),
    "flyb_peer_total_pad": (C.c_int32, [C.c_int32]),
    "flyb_peer_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "flyb_peer_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p), _P]),
    "flyb_peer_open": (C.c_int, [_P, C.POINTER(C.c_void_p)]),
    "flyb_peer_close": (C.c_int, [_P]),
    "flyb_peer_free": (C.c_int, [_P]),
    "flyb_peer_stamps": (C.c_int, [C.POINTER(PeerGroup), C.POINTER(C.c_uint64)]),
    "flyb_peer_status": (C.c_int, [C.POINTER(PeerGroup), C.POINTER(C.c_int32)]),
    "flyb_anneal_var": (C.c_int, [_P, C.c_int32, C.c_float, C.c_float, _P, _P]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libflyb200.so once.  Raises (never falls back) when it is missing or lacks a declared symbol."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FlybError(
                f"{LIB_PATH} not found: build it with `python -m fly_bproject_b200.build` "
                "(__graft_entry__.build()).  flyb200 has no CPU or PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so does not export it
            fn.restype = restype
            fn.argtypes = argtypes
        if handle.flyb_abi_version() != 7:
            raise FlybError("libflyb200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().flyb_last_error().decode("utf-8", "replace")
        raise FlybError(f"{what or 'flyb200'} failed ({rc}): {msg}")


def ptr(t) -> int | None:
    """Device pointer of a torch tensor (None -> NULL).  Tensors must be contiguous CUDA tensors."""
    if t is None:
        return None
    if not t.is_cuda:
        raise FlybError("flyb200 kernels take CUDA tensors only (no CPU fallback); got a CPU tensor")
    if not t.is_contiguous():
        raise FlybError("flyb200 kernels take contiguous tensors")
    return t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
