"""A/B of CallFrontend scheduling knobs on ONE box: e2e ms/step from pinned float32 host calls (cfg2).

Findings on the B200 box (11th session; the knobs beyond kMinGroup / kFirstGroup were experiments and are not in the
library): tapering the final sub-batches to 2, 1, 1 calls 24.65 vs 24.27-24.82 ms (noise); two copy streams 29.5 ms and
three 32.5 ms against 24.8 ms with one (although two streams ALONE copy at 54.8 GB/s against 49.9, tools/h2d_probe.py);
copies split into 8 MB / 4 MB pieces 25.2 / 25.7 ms against 24.7 ms whole.  With the kernels running, the link delivers
~49.8 GB/s whatever the copy pattern: the e2e step sits on that floor (1.2288 GB -> 24.7 ms)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth
from audio_processor_b200.frontend import CallFrontend

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2).pin_memory() for i in range(8)]
calls = [pool[i % 8] for i in range(64)]
def run(steps=10):
    cf = CallFrontend(dev)
    for _ in range(2):
        cf.process_calls(calls, 8000, acoustic_to_host=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = cf.process_calls(calls, 8000, acoustic_to_host=True)
        float(r.acoustic_rows[:, 0].sum())
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / steps


variants = {"8 MB pieces": dict(), "whole calls": dict(kCopyPieceBytes=1 << 30), "4 MB pieces": dict(kCopyPieceBytes=4 << 20),
            "10 MB pieces": dict(kCopyPieceBytes=10 << 20), "8 MB no-taper": dict(kTaper=False)}
base = dict(kTaper=True, kMinGroup=4, kFirstGroup=2, kCopyStreams=1, kCopyPieceBytes=8 << 20)
for rnd in range(3):
    for name, kw in variants.items():
        for k, v in {**base, **kw}.items():
            setattr(CallFrontend, k, v)
        print(f"round {rnd} {name}: {run():.2f} ms/step", flush=True)
