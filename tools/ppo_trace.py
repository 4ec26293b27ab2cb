#!/usr/bin/env python
"""clock64 timeline of CTA 0 of the PPO forward/backward chain kernel (debug aid; flyb_debug_set_trace).

MMA warp marks come in pairs per phase: operand seen / MMAs issued + committed.  Epilogue marks (thread 0): every
accumulator-ready wait passed and every operand hand-over, in program order.
"""
import ctypes as C
import os
import sys

if len(sys.argv) > 1 and sys.argv[1] == "dw":
    os.environ["FLYB_TRACE_DW"] = "1"      # read by the library at its first PPO gradient call

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import fly_bproject_b200 as fb  # noqa: E402
from fly_bproject_b200 import _lib  # noqa: E402

dev = "cuda:0"
agent = fb.PPO(fb.make_args(num_envs=16384, sim_device=dev, horizon=32, verbose=False, state_pool=2, seed=1))
agent.collect_rollout(use_graph=False)
agent.make_data()
for _ in range(3):
    agent._minibatch(0, agent.mini_chunk_size)
torch.cuda.synchronize()
trace = torch.zeros(192 + 2 * 148, dtype=torch.int64, device=dev)
_lib.check(agent.lib.flyb_debug_set_trace(trace.data_ptr()))
agent._minibatch(0, agent.mini_chunk_size)
torch.cuda.synchronize()
agent.lib.flyb_debug_set_trace(None)
if len(sys.argv) > 1 and sys.argv[1] == "dw":
    # the weight-gradient kernel ran last and overwrote the buffer: its marks (see DwArgs::trace)
    t = trace.cpu().tolist()
    t0 = min(x for x in t if x > 0)
    prod = [x - t0 for x in t[:64] if x > 0]
    tr = [x - t0 if x > 0 else -1 for x in t[64:127]]
    mma = [x - t0 for x in t[128:190] if x > 0]
    print("producer: raw slot free seen, per stage:", prod)
    print("transposer w0 (raw full seen, op slot free seen, handed on) per stage:",
          [tuple(tr[3 * i:3 * i + 3]) for i in range(21)])
    print("MMA warp: operands seen, per stage:", mma)
    d = [b - a for a, b in zip(mma, mma[1:])]
    print("MMA stage period (cycles): min %d median %d max %d; %d stages, last at %d" % (
        min(d), sorted(d)[len(d) // 2], max(d), len(mma), mma[-1]))
    sys.exit(0)
t = trace.cpu().tolist()
t0 = min(x for x in t[:128] if x > 0)
mm = [x - t0 for x in t[:64] if x > 0]
ep = [x - t0 for x in t[64:128] if x > 0]
print("MMA warp   :", mm)
print("epilogue t0:", ep)
names = ["X seen", "D2 issued", "A3 seen", "D3 issued", "A4 seen", "D4 issued", "A5 seen", "D5 issued", "A6 seen",
         "D6 issued", "A7 seen", "D7 issued"]
per = len(names)
for k in range(len(mm) // per):
    row = mm[k * per:(k + 1) * per]
    print(f"tile {k}: " + ", ".join(f"{n} {v}" for n, v in zip(names, row)))
    print("   waits for operands (cycles): " + ", ".join(
        f"{names[i]} {row[i] - row[i - 1]}" for i in range(2, per, 2)) + f" | tile span {row[-1] - row[0]}")
