"""Data-parallel plumbing: one process per GPU, contiguous env shards, ONE collective.

Envs are independent in every hot-path function, so rank r simply owns envs ``[r*N/G, (r+1)*N/G)`` with its own state
tensors, rollout buffers and RNG stream; rollout and GAE never communicate.  The only exchange is the all-reduce(SUM)
of the flat gradient between ``flyb_ppo_grad`` and ``flyb_ppo_apply``; the mean is taken by ``grad_scale = 1/G`` inside
``flyb_ppo_apply`` BEFORE the global-norm clip, so every rank clips and steps identically.  Equal shard sizes make the
mean of the shard means the global mean of both loss terms (reference ppo.py:194-197), i.e. a G-shard step equals the
reference's single-process step on the concatenated batch up to fp reassociation.

On NVLink-connected GPUs the exchange does not go through NCCL at all: `PeerExchange` maps every rank's exchange
buffer into every process (CUDA IPC) and the library's own kernels do the sum — inside ``flyb_ppo_step_dp`` the threads
of the gradient-reducing kernel push their entries into all ranks' buffers as self-validating 64-bit words, await the
peers' ones and add them in rank order (the split form: the reducing kernel of ``flyb_ppo_grad`` pushes, the norm kernel
of ``flyb_ppo_apply`` waits for flags and adds the slots; include/flyb200.h, FlybPeerGroup).  ``FLYB_DP=nccl`` keeps the NCCL all-reduce (the baseline the fused path is
measured against); torch.distributed is then only the side channel for the IPC handles and the parameter broadcast.

The plain helpers are device-agnostic (NCCL on the GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
from typing import List, Optional, Tuple

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


def exchange_mode() -> str:
    """'peer' (own kernels over NVLink peer memory, default) or 'nccl' (FLYB_DP=nccl)."""
    mode = os.environ.get("FLYB_DP", "peer").lower()
    if mode not in ("peer", "nccl"):
        raise ValueError(f"FLYB_DP={mode!r}: expected 'peer' or 'nccl'")
    return mode


class PeerExchange:
    """Owns this rank's exchange buffer and the mappings of the other ranks' ones; `.group` is the FlybPeerGroup the
    C ABI takes.  Collective: every rank of the default process group must construct it at the same point."""

    def __init__(self, total: int, device: torch.device):
        """Never skips a collective: a local failure is remembered in `self.error` and raised by `make_exchange`
        only after every rank has been through the same calls."""
        from . import _lib
        self._lib, self._L = _lib, _lib.lib()
        self.rank, self.world_size = world()
        self._own: Optional[int] = None
        self._opened: List[int] = []
        self.error = ""
        self.group = _lib.PeerGroup(self.rank, self.world_size, self._L.flyb_peer_total_pad(total), 0)
        raw: Optional[bytes] = None
        device = torch.device(device)
        ctx = torch.cuda.device(device) if device.type == "cuda" else contextlib.nullcontext()
        with ctx:
            try:
                if self.world_size > _lib.MAX_PEERS:
                    raise _lib.FlybError(f"covers up to {_lib.MAX_PEERS} ranks of one NVLink domain, got {self.world_size}")
                own, handle = C.c_void_p(), C.create_string_buffer(64)
                _lib.check(self._L.flyb_peer_alloc(self._L.flyb_peer_bytes(total, self.world_size), C.byref(own), handle),
                           "flyb_peer_alloc")
                self._own, raw = own.value, handle.raw
            except Exception as e:      # noqa: BLE001
                self.error = f"{type(e).__name__}: {e}"
            handles: List[Optional[bytes]] = [None] * self.world_size
            dist.all_gather_object(handles, raw)
            if not self.error and any(h is None for h in handles):
                self.error = "another rank could not allocate its exchange buffer"
            if not self.error:
                try:
                    for r in range(self.world_size):
                        if r == self.rank:
                            self.group.base[r] = self._own
                            continue
                        mapped = C.c_void_p()
                        _lib.check(self._L.flyb_peer_open(handles[r], C.byref(mapped)), f"flyb_peer_open(rank {r})")
                        self._opened.append(mapped.value)
                        self.group.base[r] = mapped.value
                except Exception as e:  # noqa: BLE001
                    self.error = f"{type(e).__name__}: {e}"

    def byref(self):
        return C.byref(self.group)

    def status(self) -> int:
        """Sticky error word of this rank's exchange (flyb_peer_status; synchronises): 0 = fine, 1 + r = the waiting kernel
        gave up waiting for rank r."""
        st = C.c_int32(0)
        self._lib.check(self._L.flyb_peer_status(self.byref(), C.byref(st)), "flyb_peer_status")
        return int(st.value)

    def check(self) -> None:
        """Raise when a wait for a peer timed out (the device did not trap; parameters are meaningless from there on)."""
        st = self.status() if self._own is not None else 0
        if st:
            raise self._lib.FlybError(
                f"rank {self.rank}: the peer-memory gradient exchange gave up waiting for rank {st - 1} "
                "(FLYB_PEER_TIMEOUT_S, default 120 s); restart from the last checkpoint, with FLYB_DP=nccl if a rank "
                "legitimately stalls that long")

    def unmap(self) -> None:
        for m in self._opened:
            self._L.flyb_peer_close(m)
        self._opened = []

    def free(self) -> None:
        if self._own is not None:
            self._L.flyb_peer_free(self._own)
            self._own = None

    def close(self) -> None:
        """Collective: unmap the peers' buffers, then (after a barrier: nobody frees what is still mapped) free the
        own one."""
        if self._own is None:
            return
        torch.cuda.synchronize()
        self.unmap()
        if dist.is_initialized():
            dist.barrier()
        self.free()


def make_exchange(total: int, device: torch.device) -> Optional[PeerExchange]:
    """Collective.  The peer exchange when every rank could map every buffer, else None (-> NCCL all-reduce) on ALL
    ranks: a rank that failed (GPUs without peer access, IPC refused by the container) must not leave the others in a
    different mode."""
    if exchange_mode() != "peer":
        return None
    ex = PeerExchange(total, device)
    ok = torch.tensor([0 if ex.error else 1], dtype=torch.int32, device=device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)      # also the barrier before the first push: all buffers zeroed + mapped
    if int(ok.item()) == 1:
        return ex
    if ex.error:
        print(f"flyb200: rank {ex.rank}: peer exchange unavailable ({ex.error}); the gradient sum uses NCCL", flush=True)
    ex.unmap()
    dist.barrier()
    ex.free()
    return None
