"""Kernel-level timing of the front end stages with CUDA events (development aid).

    python tools/microbench.py [--chunks 96] [--reps 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from audio_processor_b200 import _native, runtime  # noqa: E402


FE = None


def timed(fn, reps, flush):
    """(median ms, best ms) of the GPU work of fn(): sum of the per-kernel CUDA-event times
    recorded inside the C ABI (host-side pointer preparation is excluded)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        flush.zero_()                      # evict L2 (buffer > 126 MB)
        FE.profile(True)
        fn()
        torch.cuda.synchronize()
        prof = FE.profile_read()
        FE.profile(False)
        ms.append(sum(v[0] for v in prof.values()))
        timed.last = {k: round(v[0], 4) for k, v in prof.items() if v[1]}
    ms.sort()
    return ms[len(ms) // 2], ms[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chunks", type=int, default=96)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    global FE
    fe = FE = runtime.get_frontend(0)
    dev = fe.device
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    res = {"num_sms": fe.num_sms}
    # resample: one 5-min stereo 8 kHz call
    if not a.only or a.only == "resample":
        call = (torch.rand(2, 2_400_000, generator=g) - 0.5).to(dev)
        med, best = timed(lambda: fe.resample(call, 8000, 16000, want_mixed=True), a.reps, flush)
        byts = 2_400_000 * 2 * 4 + 3 * 4_800_000 * 4
        res["resample_8k_stereo_5min"] = {"ms_med": med, "ms_best": best, "GBps_best": byts / best / 1e6}
    if not a.only or a.only == "resample":
        calls16 = [(torch.rand(2, 2_400_000, generator=g) - 0.5).to(dev) for _ in range(16)]
        med, best = timed(lambda: fe.resample_batch(calls16, 8000, 16000, want_mixed=True), a.reps, flush)
        res["resample_8k_stereo_5min_x16_one_launch"] = {"ms_med": med, "ms_best": best, "GBps_best": 16 * byts / best / 1e6}
        del calls16
    # fbank on N full 25 s chunks
    # the bench's kind of data: speech-like 8 kHz calls (tones + noise + silent gaps) upsampled to
    # 16 kHz, cut into 25 s chunks -- timing and profiles are data-dependent (rolloff scan, floors)
    from audio_processor_b200 import synth
    n_calls = (a.chunks + 11) // 12
    src = [synth.make_call(300.0, 8000, 500 + i, 1) for i in range(min(n_calls, 4))]
    up = [fe.resample(c.to(dev), 8000, 16000) for c in src]
    chunks = [up[(i // 12) % len(up)][0, (i % 12) * 384_000:(i % 12) * 384_000 + 400_000] for i in range(a.chunks)]
    frames = a.chunks * 2498
    for name, prec in (("fbank_f64", _native.PRECISION_F64), ("fbank_f32", _native.PRECISION_F32)):
        if a.only and a.only != name:
            continue
        med, best = timed(lambda: fe.fbank(chunks, precision=prec, group_size=16), a.reps, flush)
        res[name] = {"ms_med": med, "ms_best": best, "Mframes_per_s": frames / best / 1e3,
                     "algorithmic_GBps": frames * 962 / best / 1e6, "kernels": timed.last}
        med, best = timed(lambda: fe.fbank(chunks, precision=prec, group_size=16, raw=True), a.reps, flush)
        res[name + "_rawonly"] = {"ms_med": med, "ms_best": best, "Mframes_per_s": frames / best / 1e3,
                                  "algorithmic_GBps": frames * 962 / best / 1e6}
    if not a.only or a.only == "acoustic":
        med, best = timed(lambda: fe.acoustic(chunks), a.reps, flush)
        fr2 = a.chunks * 2501
        res["acoustic"] = {"ms_med": med, "ms_best": best, "Mframes_per_s": fr2 / best / 1e3, "kernels": timed.last}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
