#!/usr/bin/env python
"""clock64 timeline of CTA 0 of the fused env-step kernel (debug aid; flyb_debug_set_trace): env phase marks, MMA warp,
epilogue warp 0 and the act finaliser, all relative to the kernel's entry.  python tools/step_trace.py [num_envs]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import fly_bproject_b200 as fb  # noqa: E402
from fly_bproject_b200 import _lib  # noqa: E402

dev = "cuda:0"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
args = fb.make_args(num_envs=n, sim_device=dev, horizon=8, verbose=False, state_pool=8, seed=1)
agent = fb.PPO(args)
assert agent._fused_steps
for _ in range(2):
    agent.collect_rollout(use_graph=False)
torch.cuda.synchronize()
for rep in range(3):
    trace = torch.zeros(192 + 2 * 148, dtype=torch.int64, device=dev)
    _lib.check(agent.lib.flyb_debug_set_trace(trace.data_ptr()))
    agent._fused_step(1 + rep)
    torch.cuda.synchronize()
    agent.lib.flyb_debug_set_trace(None)
    t = trace.cpu().tolist()
    t0 = t[190]
    print(f"--- launch {rep}: kernel entry -> exit {t[191] - t0} cycles (thread 0 of CTA 0)")
    print("env phase [entered, warps met, tables built, root+scalars issued, all loads issued, tile staged, math done, reward+masks done]:", [x - t0 for x in t[172:190] if x > 0])
    print("MMA warp   :", [x - t0 for x in t[:64] if x > 0])
    print("epilogue w0:", [x - t0 for x in t[64:128] if x > 0])
    print("finaliser  :", [x - t0 for x in t[128:160] if x > 0])
    ins = [t[192 + 2 * b] for b in range(148) if t[192 + 2 * b] > 0]
    outs = [t[193 + 2 * b] for b in range(148) if t[193 + 2 * b] > 0]
    if ins:
        g0 = min(ins)
        srt = lambda v: sorted(x - g0 for x in v)
        si, so = srt(ins), srt(outs)
        q = lambda v, f: v[min(len(v) - 1, int(f * len(v)))]
        print(f"per-CTA global timer, ns after the first CTA's entry ({len(ins)} CTAs): entry min/med/max "
              f"{si[0]}/{q(si, .5)}/{si[-1]}   exit min/med/p90/max {so[0]}/{q(so, .5)}/{q(so, .9)}/{so[-1]}   "
              f"CTA 0: entry {t[192] - g0} exit {t[193] - g0}")
        durs = sorted(t[193 + 2 * b] - t[192 + 2 * b] for b in range(148) if t[192 + 2 * b] > 0)
        print(f"per-CTA duration ns min/med/p90/max {durs[0]}/{q(durs, .5)}/{q(durs, .9)}/{durs[-1]}")
