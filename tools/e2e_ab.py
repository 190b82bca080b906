"""A/B of CallFrontend scheduling knobs on ONE box: e2e ms/step from pinned host calls (cfg2), float32 and 16-bit wav.

Findings of the 11th session (knobs that were experiments only are not in the library): two copy streams 29.5 ms and three
32.5 ms against 24.8 ms with one (although two streams ALONE copy at 54.8 GB/s against 49.9, tools/h2d_probe.py); copies split
into 8 MB / 4 MB pieces 25.2 / 25.7 ms against 24.7 ms whole; tapering the last sub-batches: noise; with the group's copies
waited for on every copy stream (second try): two streams 27.2 ms, three 30.5 ms against 24.2 ms; copies enqueued 16
calls ahead of the launches instead of all 64 first: 17.16 vs 17.30 ms on the wav path, nothing on float32 -- not kept.  tools/e2e_timeline.py
shows why: with 4-call sub-batches the GPU is busy almost all the time (1.2-1.9 ms per 4 calls against 1.55 ms of upload)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth, wav
from audio_processor_b200.frontend import CallFrontend

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(8)]
pin = [c.pin_memory() for c in pool]
imgs = [synth.wav_image(c, 8000).pin_memory() for c in pool]
inputs = {"f32": lambda: [pin[i % 8] for i in range(64)],
          "wav16": lambda: [wav.load_pcm16(imgs[i % 8])[0] for i in range(64)]}


def run(kind, steps=10):
    cf = CallFrontend(dev)
    for _ in range(2):
        cf.process_calls(inputs[kind](), 8000, acoustic_to_host=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = cf.process_calls(inputs[kind](), 8000, acoustic_to_host=True)
        float(r.acoustic_rows[:, 0].sum())
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / steps


variants = {"two streams": dict(), "one stream": dict(kTwoStreams=False)}
base = dict(kMinGroup=4, kFirstGroup=2, kMaxGroup=16, kTwoStreams=True)
for rnd in range(3):
    for name, kw in variants.items():
        for k, v in {**base, **kw}.items():
            setattr(CallFrontend, k, v)
        print(f"round {rnd} {name}: f32 {run('f32'):.2f}  wav16 {run('wav16'):.2f} ms/step", flush=True)
