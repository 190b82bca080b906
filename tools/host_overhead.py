"""Host-side cost of CallFrontend.process_calls: time to ENQUEUE a step vs time for the GPU to finish it."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth
from audio_processor_b200.frontend import CallFrontend

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2).pin_memory() for i in range(8)]
host_calls = [pool[i % 8] for i in range(64)]
dev_calls = [c.to(dev) for c in pool]
dev_calls = [dev_calls[i % 8] for i in range(64)]
cf = CallFrontend(dev)
for name, calls, kw in (("device", dev_calls, dict(acoustic_to_host=False)), ("host sb8", host_calls, dict(acoustic_to_host=False, sub_batch=8)),
                        ("host sb4", host_calls, dict(acoustic_to_host=False, sub_batch=4))):
    for _ in range(3):
        r = cf.process_calls(calls, 8000, **kw)
    torch.cuda.synchronize()
    enq, tot = [], []
    for _ in range(8):
        t0 = time.perf_counter()
        r = cf.process_calls(calls, 8000, **kw)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        enq.append(1e3 * (t1 - t0)); tot.append(1e3 * (t2 - t0))
    print(f"{name}: enqueue {sorted(enq)[len(enq)//2]:.2f} ms, total {sorted(tot)[len(tot)//2]:.2f} ms")

# kernel time inside a pipelined host step (sum of per-kernel CUDA-event times) vs its wall time
fe = cf.fe
for sbv in (4, 8):
    fe.profile(True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        r = cf.process_calls(host_calls, 8000, acoustic_to_host=True, sub_batch=sbv)
    torch.cuda.synchronize()
    wall = 1e3 * (time.perf_counter() - t0) / 5
    prof = fe.profile_read()
    fe.profile(False)
    ksum = sum(v[0] for v in prof.values()) / 5
    print(f"host sb{sbv}: wall {wall:.2f} ms/step, kernels {ksum:.2f} ms/step,", {k: round(v[0] / 5, 2) for k, v in prof.items() if v[1]})
