"""A/B timing of kernel variants (development aid; one gpurun call times many builds).

    python tools/kernel_lab.py build  name1:-DAFE_X=1,-DAFE_Y=2  name2:  ...     # here (no GPU): tools/_bin/libafe_<name>.so
    python tools/kernel_lab.py run [--family fbank_f64|acoustic] [--chunks 96] [--reps 7] [--trace]   # on the GPU box

`run` times every tools/_bin/libafe_*.so on the microbench batch (per-kernel CUDA events recorded inside the
C ABI), checks every variant's output against the first one, and prints one line per variant.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, "tools", "_bin")


def build(specs):
    from audio_processor_b200.csrc import build as b
    os.makedirs(BIN, exist_ok=True)
    for spec in specs:
        name, _, flags = spec.partition(":")
        extra = [f for f in flags.split(",") if f]
        out = os.path.join(BIN, f"libafe_{name}.so")
        b.build_cuda(force=True, extra=extra, out=out)
        print("built", out, extra)


def make_frontend(path):
    import torch
    from audio_processor_b200 import _native, runtime
    fe = runtime.Frontend.__new__(runtime.Frontend)
    fe.lib = _native.Library(path)
    fe.device = torch.device("cuda", 0)
    fe.ctx = fe.lib.create(0)
    fe.num_sms = int(fe.lib.afe_num_sms(fe.ctx))
    fe._tls = threading.local()
    return fe


def run(a):
    import torch
    from audio_processor_b200 import _native, synth
    libs = sorted(glob.glob(os.path.join(BIN, "libafe_*.so")))
    if a.only:
        libs = [l for l in libs if any(o in os.path.basename(l) for o in a.only.split(","))]
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ref = None
    first = make_frontend(libs[0])
    n_calls = (a.chunks + 11) // 12
    src = [synth.make_call(300.0, 8000, 500 + i, 1) for i in range(min(n_calls, 4))]
    up = [first.resample(c.to(dev), 8000, 16000) for c in src]
    chunks = [up[(i // 12) % len(up)][0, (i % 12) * 384_000:(i % 12) * 384_000 + 400_000] for i in range(a.chunks)]
    for path in libs:
        name = os.path.basename(path)[len("libafe_"):-3]
        fe = make_frontend(path)
        if a.family.startswith("fbank"):
            prec = _native.PRECISION_F64 if a.family == "fbank_f64" else _native.PRECISION_F32
            fn = lambda: fe.fbank(chunks, precision=prec, group_size=16)
            flat = lambda o: torch.cat([f.reshape(-1) for f, _ in o])
        else:
            fn = lambda: fe.acoustic(chunks)
            flat = lambda o: o[0].reshape(-1)
        for _ in range(3):
            out = fn()
        torch.cuda.synchronize()
        vec = flat(out).double()
        if ref is None:
            ref = vec
            err = 0.0
        else:
            err = float((vec - ref).abs().max())
        tot, kern = [], {}
        for _ in range(a.reps):
            flush.zero_()
            fe.profile(True)
            fn()
            torch.cuda.synchronize()
            prof = fe.profile_read()
            fe.profile(False)
            tot.append(sum(v[0] for v in prof.values()))
            kern = {k: round(v[0], 4) for k, v in prof.items() if v[1]}
        tot.sort()
        print(json.dumps({"variant": name, "ms_med": round(tot[len(tot) // 2], 4), "ms_best": round(tot[0], 4),
                          "kernels": kern, "max_abs_vs_first": err}), flush=True)
        if a.trace:
            os.environ["AFE_PHASE_TRACE"] = "1"
            sub = subprocess.run([sys.executable, __file__, "run", "--family", a.family, "--chunks", str(a.chunks),
                                  "--reps", "1", "--only", os.path.basename(path)], capture_output=True, text=True,
                                 env={**os.environ, "AFE_PHASE_TRACE": "1"})
            del os.environ["AFE_PHASE_TRACE"]
            lines = [l for l in sub.stderr.splitlines() if "afe trace" in l]
            for l in lines[-17:]:
                print("   ", l[:220])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        ap = argparse.ArgumentParser()
        ap.add_argument("cmd")
        ap.add_argument("--family", default="fbank_f64")
        ap.add_argument("--chunks", type=int, default=96)
        ap.add_argument("--reps", type=int, default=7)
        ap.add_argument("--only", default="")
        ap.add_argument("--trace", action="store_true")
        run(ap.parse_args())
