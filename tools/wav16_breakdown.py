"""Where the e2e_wav16 step goes: header walk on the host, enqueue, kernels, wall (cfg2, 64 calls)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from audio_processor_b200 import synth, wav
from audio_processor_b200.frontend import CallFrontend

dev = torch.device("cuda", 0)
pool = [synth.make_call(300.0, 8000, 100 + i, 2) for i in range(8)]
wav_pool = [synth.wav_image(c, 8000).pin_memory() for c in pool]
files = [wav_pool[i % 8] for i in range(64)]
cf = CallFrontend(dev); fe = cf.fe
for sb in (None, 4, 8, 16, 64):
    for _ in range(2):
        cf.process_calls([wav.load_pcm16(f)[0] for f in files], 8000, sub_batch=sb)
    torch.cuda.synchronize()
    tp = te = tw = 0.0
    fe.profile(True)
    n = 6
    for _ in range(n):
        t0 = time.perf_counter()
        pcm = [wav.load_pcm16(f)[0] for f in files]
        t1 = time.perf_counter()
        r = cf.process_calls(pcm, 8000, acoustic_to_host=True, sub_batch=sb)
        t2 = time.perf_counter()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        tp += t1 - t0; te += t2 - t1; tw += t3 - t0
    prof = fe.profile_read(); fe.profile(False)
    ks = {k: round(v[0] / n, 2) for k, v in prof.items() if v[1]}
    nl = sum(v[1] for v in prof.values()) / n
    print(f"sub_batch {sb}: parse {1e3*tp/n:.2f} ms, process_calls {1e3*te/n:.2f} ms, wall {1e3*tw/n:.2f} ms, kernels {sum(ks.values()):.2f} ms in {nl:.0f} launches {ks}", flush=True)
