"""Data-parallel plumbing: one process per GPU, contiguous env shards, ONE collective.

Envs are independent in every hot-path function, so rank r simply owns envs ``[r*N/G, (r+1)*N/G)`` with its own state
tensors, rollout buffers and RNG stream; rollout and GAE never communicate.  The only exchange is the all-reduce(SUM)
of the flat gradient between ``flyb_ppo_grad`` and ``flyb_ppo_apply``; the mean is taken by ``grad_scale = 1/G`` inside
``flyb_ppo_apply`` BEFORE the global-norm clip, so every rank clips and steps identically.  Equal shard sizes make the
mean of the shard means the global mean of both loss terms (reference ppo.py:194-197), i.e. a G-shard step equals the
reference's single-process step on the concatenated batch up to fp reassociation.

These helpers are device-agnostic (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    """(rank, world_size); (0, 1) outside a process group."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def env_shard(num_envs_total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """(first env, env count) of `rank`'s contiguous block.  Shards must be equal (see module docstring)."""
    if world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if num_envs_total % world_size != 0:
        raise ValueError(f"num_envs_total={num_envs_total} is not divisible by world_size={world_size}: unequal "
                         "shards would make the mean of shard means differ from the global mean")
    per = num_envs_total // world_size
    return rank * per, per


def grad_all_reduce_(flat_grad: torch.Tensor) -> float:
    """In-place all-reduce(SUM) of the flat gradient; returns the `grad_scale` (1/G) flyb_ppo_apply must use."""
    _, g = world()
    if g > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM)
    return 1.0 / g


def broadcast_params_(flat_params: torch.Tensor, src: int = 0) -> None:
    _, g = world()
    if g > 1:
        dist.broadcast(flat_params, src=src)


def rank_seed(seed: int) -> int:
    """RNG stream of this rank (seed + rank)."""
    return int(seed) + world()[0]
