"""TEST INFRASTRUCTURE -- generate tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python -m oracle.make_golden

Every array stored here is an input to, or an output of, one of the reference's own
hot-path call sites executed on the CPU with the installed
torch / torchaudio / transformers.  The fixtures travel to the GPU box; the reference
does not.
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from audio_processor_b200 import synth  # noqa: E402
from oracle import reference_calls as rc  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def gen_resample(ap):
    out = {}
    # 8 kHz stereo, 1.5 s: resample + split through split_audio itself
    x8 = synth.make_call(1.5, 8000, 11, 2)
    chunks, agent, client = ap.split_audio(x8.clone(), 8000, "g.wav")
    out["x8"] = x8.numpy()
    out["agent16"] = agent.numpy()
    out["client16"] = client.numpy()
    out["mixed16"] = chunks[0]["stereo_waveform"].numpy()
    # white-noise variant, odd length
    xw = torch.stack([synth.white(4001, 21), synth.white(4001, 22)], 0)
    chunks, agent, client = ap.split_audio(xw.clone(), 8000, "w.wav")
    out["xw8"] = xw.numpy()
    out["w_agent16"] = agent.numpy(); out["w_client16"] = client.numpy()
    out["w_mixed16"] = chunks[0]["stereo_waveform"].numpy()
    # 44.1 kHz mono, 0.3 s (13230 samples) through get_resampler
    x44 = synth.make_call(0.3, 44100, 5, 1)
    out["x44"] = x44.numpy()
    out["y44to16"] = ap.get_resampler(44100, 16000)(x44).numpy()
    # >= 1 s: torch's conv1d takes the in-order FMA path and the oracle is bit-exact; shorter
    # inputs take another oneDNN blocking and differ from it by <= 1 ulp on ~2.5 % of samples
    x44l = synth.make_call(1.0, 44100, 6, 1)
    out["x44_long"] = x44l.numpy()
    out["y44to16_long"] = ap.get_resampler(44100, 16000)(x44l).numpy()
    # tap banks
    out["taps_8_16"] = ap.get_resampler(8000, 16000).kernel.numpy()[:, 0, :]
    out["taps_441_160"] = ap.get_resampler(44100, 16000).kernel.numpy()[:, 0, :]
    # mono 8 kHz: all three outputs alias
    xm = synth.make_call(0.5, 8000, 31, 1)
    chunks, agent, client = ap.split_audio(xm.clone(), 8000, "m.wav")
    out["xm8"] = xm.numpy(); out["m_agent16"] = agent.numpy()
    return out


def gen_fbank(ap, proc):
    out = {}
    lengths = [399, 400, 559, 560, 720, 5000, 16000]
    waves = [synth.speech_like(n, 16000, 40 + i).numpy() for i, n in enumerate(lengths)]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        res = proc(waves, sampling_rate=16000, return_tensors="pt", padding=True)
    out["ragged_lengths"] = np.asarray(lengths, np.int64)
    out["ragged_waves"] = np.concatenate(waves)
    out["ragged_features"] = res["input_features"].numpy()
    out["ragged_mask"] = res["attention_mask"].numpy()
    # single-item calls at the degenerate lengths (speech_segment.py:264 path)
    for n in (399, 400, 559, 560, 720):
        w = synth.speech_like(n, 16000, 90 + n).numpy()
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            r = proc(w, sampling_rate=16000, return_tensors="pt", padding=True)
        out[f"single_{n}_wave"] = w
        out[f"single_{n}_features"] = r["input_features"].numpy()
        out[f"single_{n}_mask"] = r["attention_mask"].numpy()
    # 8 kHz call resampled by the reference, 3 s: the empty 4-8 kHz band case
    x8 = synth.make_call(3.0, 8000, 12, 2)
    _, agent, _ = ap.split_audio(x8.clone(), 8000, "u.wav")
    up = agent[0].numpy()
    white = synth.white(32000, 3).numpy()
    tone = (0.4 * np.sin(2 * np.pi * 1000.0 * np.arange(16000) / 16000.0)).astype(np.float32)
    zeros = np.zeros(4000, np.float32)
    for name, w in (("up8k", up), ("white", white), ("tone1k", tone), ("zeros", zeros)):
        r = proc([w], sampling_rate=16000, return_tensors="pt", padding=True)
        out[f"{name}_wave"] = w
        out[f"{name}_features"] = r["input_features"].numpy()
        out[f"{name}_mask"] = r["attention_mask"].numpy()
        out[f"{name}_raw"] = proc._extract_fbank_features(w)
    out["mel_filters"] = proc.mel_filters
    out["window"] = proc.window
    return out


def gen_acoustic(an):
    out = {}
    cases = {}
    for n in (400, 479, 560, 8000, 16002, 48000):
        cases[f"speech_{n}"] = synth.speech_like(n, 16000, 60 + n % 97).numpy()
    cases["white_32000"] = synth.white(32000, 8).numpy()
    cases["zeros_16000"] = np.zeros(16000, np.float32)
    cases["tone_16000"] = (0.4 * np.sin(2 * np.pi * 1000.0 * np.arange(16000) / 16000.0)).astype(np.float32)
    sq = np.where((np.arange(16000) // 20) % 2 == 0, 0.5, -0.5).astype(np.float32)
    cases["square_16000"] = sq
    cases["dc_16000"] = np.full(16000, 0.25, np.float32)
    x8 = synth.make_call(2.0, 8000, 13, 2)
    cases["up8k_32000"] = rc.audio_processor().split_audio(x8.clone(), 8000, "a.wav")[2][0].numpy()
    names = []
    for name, w in cases.items():
        f = an._extract_acoustic_features(torch.from_numpy(w)[None, :], 16000)
        names.append(name)
        out[f"{name}_wave"] = w
        out[f"{name}_values"] = np.asarray(list(f.values()), np.float64)
        out[f"{name}_keys"] = np.asarray(list(f.keys()))
    out["names"] = np.asarray(names)
    out["zero_keys"] = np.asarray(list(an._get_zero_features().keys()))
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        short = an._extract_acoustic_features(torch.zeros(1, 399), 16000)
    out["short_is_none"] = np.asarray([short is None])
    return out


def gen_tables(ap, seg):
    out = {}
    totals = [0, 1, 399999, 400000, 400001, 784000, 784001, 960000, 4800000]
    rows = []
    for tot in totals:
        chunks, _, _ = ap.split_audio(torch.zeros(1, tot), 16000, "t.wav")
        for c in chunks:
            n = c["stereo_waveform"].shape[1]
            start = int(round(c["start_time"] * 16000))
            rows.append((tot, c["chunk_idx"], start, start + n, c["start_time"], c["end_time"]))
    out["chunk_rows"] = np.asarray(rows, np.float64)
    out["chunk_totals"] = np.asarray(totals, np.int64)
    # segments: reference config (25 s windows) and code default (20 s)
    wf = torch.zeros(1, 16000 * 100 + 37)
    intervals = [(0.2, 1.1), (3.337, 31.9), (40.0, 99.99), (60.123456, 60.2), (99.5, 100.5)]
    for tag, cfg in (("cfg", dict(rc.DEFAULT_CONFIG)), ("default", {k: v for k, v in rc.DEFAULT_CONFIG.items()
                                                               if not k.startswith("vad_")})):
        seg.config = cfg
        ex = seg._extract_segments(wf, 16000, intervals, "agent", 0.5)
        out[f"seg_{tag}"] = np.asarray([(e["segment_waveform"].shape[1], e["start"], e["end"]) for e in ex], np.float64)
    seg.config = dict(rc.DEFAULT_CONFIG)
    out["seg_intervals"] = np.asarray(intervals, np.float64)
    out["seg_num_samples"] = np.asarray([wf.shape[1]], np.int64)

    class T:  # the VAD timeline objects only need .start / .end
        def __init__(self, s, e): self.start, self.end = s, e
    tl = [T(0.0, 1.0), T(1.5, 2.0), T(2.81, 3.0), T(10.0, 11.0), T(11.8, 12.0)]
    out["merge_in"] = np.asarray([(t.start, t.end) for t in tl])
    out["merge_out"] = np.asarray(seg._merge_segments(tl, 0.8))
    return out


def main():
    os.makedirs(GOLD, exist_ok=True)
    ap = rc.audio_processor()
    proc = rc.processor()
    an = rc.acoustic_analyzer()
    seg = rc.speech_segment()
    import torchaudio, transformers
    meta = {"torch": torch.__version__, "torchaudio": torchaudio.__version__,
            "transformers": transformers.__version__, "numpy": np.__version__,
            "generator": "oracle/make_golden.py", "reference": rc.REFERENCE_ROOT}
    for name, fn in (("resample", lambda: gen_resample(ap)), ("fbank", lambda: gen_fbank(ap, proc)),
                     ("acoustic", lambda: gen_acoustic(an)), ("tables", lambda: gen_tables(ap, seg))):
        data = fn()
        np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **data)
        print(name, {k: getattr(v, "shape", None) for k, v in list(data.items())[:4]}, "...")
    with open(os.path.join(GOLD, "META.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
