"""Where the 16-bit wav route loses time against the device-resident step: sub-batching alone (device inputs),
then host inputs with fixed and adaptive groups.   python tools/e2e_sweep2.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from audio_processor_b200 import synth, wav  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(8)]
images = [synth.wav_image(c, 8000).pin_memory() for c in pool]
files = [images[i % 8] for i in range(64)]
pcm = [wav.load_pcm16(f)[0] for f in files]
dev_pcm = [p.to(dev) for p in pcm[:8]]
dev_calls = [dev_pcm[i % 8] for i in range(64)]
cf = CallFrontend(dev)


def run(label, calls, steps=10, **kw):
    for _ in range(3):
        r = cf.process_calls(calls, 8000, **kw)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = cf.process_calls(calls, 8000, **kw)
        s = float(r.acoustic_rows[:, 0].sum())
    torch.cuda.synchronize()
    print("%-44s %.2f ms/step" % (label, 1e3 * (time.perf_counter() - t0) / steps), flush=True)


t0 = time.perf_counter()
for _ in range(10):
    _ = [wav.load_pcm16(f)[0] for f in files]
print("64 x load_pcm16 (header walk + views): %.2f ms" % (1e2 * (time.perf_counter() - t0)), flush=True)
run("device wav frames, one group", dev_calls)
for sb in (32, 16, 8, 4):
    run("device wav frames, sub_batch %d" % sb, dev_calls, sub_batch=sb)
run("host wav frames, adaptive (2,4,16)", pcm)
for sb in (32, 16, 8, 4):
    run("host wav frames, sub_batch %d" % sb, pcm, sub_batch=sb)
for first, mn, mx in [(2, 4, 32), (4, 8, 32), (2, 8, 64), (1, 2, 8)]:
    CallFrontend.kFirstGroup, CallFrontend.kMinGroup, CallFrontend.kMaxGroup = first, mn, mx
    run("host wav frames, adaptive (%d,%d,%d)" % (first, mn, mx), pcm)
