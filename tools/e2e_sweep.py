"""Sweep the adaptive sub-batch parameters of CallFrontend.process_calls on the bench workload (e2e, float32 host input).

    python tools/e2e_sweep.py [--steps 10]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from audio_processor_b200 import synth  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--calls", type=int, default=64)
a = ap.parse_args()
dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2).pin_memory() for i in range(8)]
calls = [pool[i % 8] for i in range(a.calls)]
cf = CallFrontend(dev)
# raw PCIe rate of the same bytes, for reference
buf = torch.empty((a.calls, 2, 2_400_000), dtype=torch.float32, device=dev)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i, c in enumerate(calls):
    buf[i].copy_(c, non_blocking=True)
torch.cuda.synchronize()
t1 = time.perf_counter()
for i, c in enumerate(calls):
    buf[i].copy_(c, non_blocking=True)
torch.cuda.synchronize()
t2 = time.perf_counter()
print("plain upload of the step's calls: %.2f ms (%.1f GB/s)" % (1e3 * (t2 - t1), buf.numel() * 4 / (t2 - t1) / 1e9), flush=True)
del buf
for first, mn, mx in [(2, 4, 16), (1, 2, 16), (1, 1, 16), (1, 2, 4), (1, 1, 2), (1, 1, 1), (2, 2, 2), (2, 3, 8)]:
    CallFrontend.kFirstGroup, CallFrontend.kMinGroup, CallFrontend.kMaxGroup = first, mn, mx
    for _ in range(3):
        r = cf.process_calls(calls, 8000)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        r = cf.process_calls(calls, 8000)
        s = float(r.acoustic_rows[:, 0].sum())
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0) / a.steps
    print("first %d min %d max %d: %.2f ms/step" % (first, mn, mx, ms), flush=True)
