"""Sweep of CallFrontend's adaptive sub-batch parameters (kFirstGroup, kMinGroup, kMaxGroup) on the cfg2 step,
from pinned float32 calls and from pinned 16-bit wav images (development aid).

    python tools/e2e_group_sweep.py [--calls 64] [--steps 8]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from audio_processor_b200 import synth, wav  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402


def timed(cf, make_calls, steps):
    for _ in range(2):
        cf.process_calls(make_calls(), 8000, acoustic_to_host=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        cf.clear_plans()
        r = cf.process_calls(make_calls(), 8000, acoustic_to_host=True)
        float(r.acoustic_rows[:, 0].sum())
    e1.record()
    torch.cuda.synchronize()
    return max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - t0)) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=64)
    ap.add_argument("--steps", type=int, default=8)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    pool = [synth.make_call(300.0, 8000, 100 + i, 2).pin_memory() for i in range(a.calls)]
    images = [synth.wav_image(c, 8000).pin_memory() for c in pool]
    cf = CallFrontend(dev)
    for first, mn, mx in ((2, 4, 16), (2, 4, 32), (2, 8, 32), (4, 8, 64), (2, 16, 64), (4, 16, 32), (8, 16, 64), (2, 2, 8)):
        cf.kFirstGroup, cf.kMinGroup, cf.kMaxGroup = first, mn, mx
        f32 = timed(cf, lambda: pool, a.steps)
        w16 = timed(cf, lambda: [wav.load_pcm16(f)[0] for f in images], a.steps)
        print(json.dumps({"first": first, "min": mn, "max": mx, "e2e_f32_ms": round(f32, 2), "e2e_wav16_ms": round(w16, 2)}), flush=True)


if __name__ == "__main__":
    main()
