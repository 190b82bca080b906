"""Where the HOST time of CallFrontend.process_calls goes (development aid): cProfile of K steps on device-resident
calls, without synchronising inside the steps, layout planned every step.

    python tools/host_profile.py [--calls 64] [--steps 5] [--wav16]
"""
import argparse
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from audio_processor_b200 import synth, wav  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=64)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--wav16", action="store_true")
    ap.add_argument("--keep-plan", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(min(a.calls, 8))]
    if a.wav16:
        images = [synth.wav_image(c, 8000).pin_memory() for c in pool]
        make = lambda: [wav.load_pcm16(images[i % len(images)])[0] for i in range(a.calls)]   # noqa: E731
    else:
        devp = [c.to(dev) for c in pool]
        make = lambda: [devp[i % len(devp)] for i in range(a.calls)]                            # noqa: E731
    cf = CallFrontend(dev)
    for _ in range(2):
        cf.process_calls(make(), 8000, acoustic_to_host=False)
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    for _ in range(a.steps):
        if not a.keep_plan:
            cf.clear_plans()
        cf.process_calls(make(), 8000, acoustic_to_host=False)
    pr.disable()
    host = (time.perf_counter() - t0) / a.steps
    torch.cuda.synchronize()
    print(f"host time per step: {host * 1e3:.2f} ms ({'wav16 pinned' if a.wav16 else 'device-resident'}, plan {'cached' if a.keep_plan else 'rebuilt'})")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
