"""A/B: acoustic chain on a second stream beside fbank (CallFrontend.kTwoStreams), device-resident cfg2 step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth
from audio_processor_b200.frontend import CallFrontend
dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2).to(dev) for i in range(8)]
calls = [pool[i % 8] for i in range(64)]
for rnd in range(3):
    for two in (True, False):
        CallFrontend.kTwoStreams = two
        cf = CallFrontend(dev)
        for _ in range(3):
            cf.process_calls(calls, 8000, acoustic_to_host=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            r = cf.process_calls(calls, 8000, acoustic_to_host=False); del r
        e1.record(); torch.cuda.synchronize()
        print(f"round {rnd} two_streams={two}: {e0.elapsed_time(e1) / 10:.3f} ms/step", flush=True)
