"""Where one data-parallel optimiser step spends its exchange time (run under torchrun, one rank per GPU).

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_stamps.py

Runs a rollout and a few eager optimiser steps, then prints every rank's %globaltimer marks of the last exchange
(flyb_peer_stamps) relative to the earliest reduce-kernel entry.
"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fly_bproject_b200 as fb  # noqa: E402
from fly_bproject_b200 import _lib  # noqa: E402


def main():
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    agent = fb.PPO(fb.make_args(num_envs=16384, sim_device=f"cuda:{torch.cuda.current_device()}", horizon=32,
                                verbose=False, state_pool=4, seed=7))
    agent.collect_rollout(use_graph=False)
    agent.make_data()
    rows = []
    for it in range(12):
        agent._minibatch(0, agent.mini_chunk_size)
        out = (C.c_uint64 * 8)()
        _lib.check(agent.lib.flyb_peer_stamps(agent._peers.byref(), out), "flyb_peer_stamps")
        rows.append([int(x) for x in out])
    all_rows = [None] * dist.get_world_size()
    dist.all_gather_object(all_rows, rows)
    if rank == 0:
        if agent._dp_step == "fused":   # everything happens inside the reducing kernel
            names, cols = ["CTA 0 thread 0: quad pushed", "every rank's quad seen + summed", "CTA 0 done"], (3, 4, 5)
        else:
            names, cols = ["norm kernel entry (flags raised)", "all flags seen", "norm CTA 0 done"], (3, 4, 5)
        print(f"FLYB_DP_STEP={agent._dp_step}")
        for it in range(4, 12):     # %globaltimer is per GPU: only differences within a rank mean anything
            print(f"step {it}: " + " | ".join(
                f"r{k}: " + " ".join(f"{(r[it][j] - r[it][0]) / 1e3:6.1f}" for j in cols) for k, r in enumerate(all_rows)))
        print("columns per rank (us after that rank's reduce-kernel entry): " + ", ".join(names))
    agent.exit()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
