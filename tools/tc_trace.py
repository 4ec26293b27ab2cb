#!/usr/bin/env python
"""Print the clock64 timeline CTA 0 of the tensor-core act kernel records (debug aid; see flyb_debug_set_trace)."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import fly_bproject_b200 as fb  # noqa: E402
from fly_bproject_b200 import _lib  # noqa: E402

dev = "cuda:0"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
args = fb.make_args(num_envs=n, sim_device=dev, horizon=8, verbose=False, state_pool=2, seed=1)
agent = fb.PPO(args)
for i in range(8):
    agent._step_kernels(i)
torch.cuda.synchronize()
trace = torch.zeros(192 + 2 * 148, dtype=torch.int64, device=dev)
_lib.check(agent.lib.flyb_debug_set_trace(trace.data_ptr()))
agent._step_kernels(0)
torch.cuda.synchronize()
agent.lib.flyb_debug_set_trace(None)
t = trace.cpu().tolist()
t0 = min(x for x in t[:160] + t[190:] if x > 0)     # act kernel marks only ([160,172): post_physics)
print("MMA thread :", [x - t0 for x in t[:64] if x > 0])
print("epilogue w0:", [x - t0 for x in t[64:128] if x > 0])
print("finaliser  :", [x - t0 for x in t[128:160] if x > 0])
print("kernel entry / exit (thread 0):", t[190] - t0, t[191] - t0)
pp = t[160:172]
print("post_physics CTA 0 thread 0 (entry, loads issued, loads consumed, math done, reward+masks, mask published, obs out):",
      [x - pp[0] for x in pp[:7] if x > 0])
