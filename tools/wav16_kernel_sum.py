"""Per-kernel device time of one cfg2 step: one launch sequence over 64 device-resident calls against the adaptive
sub-batches of the 16-bit wav path (development aid; CUDA events inside the C ABI, acoustic chain on the fbank stream)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from audio_processor_b200 import synth, wav  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(64)]
images = [synth.wav_image(c, 8000).pin_memory() for c in pool]
devp = [c.to(dev) for c in pool]
cf = CallFrontend(dev)
fe = cf.fe
for name, make, two in (("resident, one launch sequence", lambda: devp, False),
                        ("wav16 pinned, adaptive sub-batches, one stream", lambda: [wav.load_pcm16(f)[0] for f in images], False),
                        ("wav16 pinned, adaptive sub-batches, two streams", lambda: [wav.load_pcm16(f)[0] for f in images], True)):
    cf.kTwoStreams = two
    for _ in range(2):
        cf.process_calls(make(), 8000, acoustic_to_host=True)
    torch.cuda.synchronize()
    steps = 5
    fe.profile(True)
    t0 = time.perf_counter()
    for _ in range(steps):
        cf.clear_plans()
        r = cf.process_calls(make(), 8000, acoustic_to_host=True)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / steps
    prof = fe.profile_read()
    fe.profile(False)
    tot = sum(v[0] for v in prof.values()) / steps
    print(f"{name}: wall {wall * 1e3:.2f} ms/step, kernels {tot:.2f} ms/step in {sum(v[1] for v in prof.values()) / steps:.0f} launches: "
          + ", ".join(f"{k} {v[0] / steps:.2f}" for k, v in prof.items() if v[1]))
