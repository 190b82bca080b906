"""CUDA-event timeline of one pipelined step (cfg2, 64 calls): when each call's upload completes and when each
sub-batch starts / ends on the compute stream, in ms from the start of the step.
    python tools/e2e_timeline.py [f32|wav16]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth, wav
from audio_processor_b200.frontend import CallFrontend

kind = sys.argv[1] if len(sys.argv) > 1 else "wav16"
dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(8)]
if kind == "wav16":
    imgs = [synth.wav_image(c, 8000).pin_memory() for c in pool]
    calls = [wav.load_pcm16(imgs[i % 8])[0] for i in range(64)]
else:
    pin = [c.pin_memory() for c in pool]
    calls = [pin[i % 8] for i in range(64)]
cf = CallFrontend(dev)
for _ in range(3):
    cf.process_calls(calls, 8000)
torch.cuda.synchronize()
for rep in range(2):
    cf.trace = {}
    r = cf.process_calls(calls, 8000)
    torch.cuda.synchronize()
    tr, cf.trace = cf.trace, None
    t0 = tr["start"]
    ups = [t0.elapsed_time(e) for e in tr["uploads"] if e is not None]
    print(f"[{kind} rep {rep}] uploads done at: " + " ".join(f"{u:.1f}" for u in ups[::4]) + f"  (last {ups[-1]:.2f})")
    busy = 0.0
    for k0, k1, g0, g1 in tr["groups"]:
        a, b = t0.elapsed_time(g0), t0.elapsed_time(g1)
        busy += b - a
        print(f"   calls {k0:2d}-{k1 - 1:2d}: {a:6.2f} -> {b:6.2f} ms ({b - a:5.2f}), last upload of the group at {ups[k1 - 1]:6.2f}")
    print(f"   span of the groups {busy:.2f} ms, step end {t0.elapsed_time(tr['groups'][-1][3]):.2f} ms")
