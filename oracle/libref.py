"""TEST / BASELINE INFRASTRUCTURE -- the reference's CPU front end, restated over the SAME
third-party libraries it calls (torchaudio Resample / MFCC, torch.stft, transformers
SeamlessM4TFeatureExtractor), because ``/root/reference`` itself cannot travel to the GPU box.

This is what ``bench.py``'s ``cpu_baseline`` leg and ``--impl reference`` time on the host
cores (``kind: "port"``).  Call sites restated: services/audio_processor.py:99-160 and
:170-181, :253-261; services/sentiment_analysis.py:807-894.  Checked against the
unmodified reference in tests/test_oracle_vs_reference.py (container only).
"""
from __future__ import annotations

import functools
import math
import os
import time
import warnings
from collections import OrderedDict
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from .np_acoustic import FEATURE_KEYS


@functools.lru_cache(maxsize=16)
def _resampler(orig: int, new: int):
    import torchaudio
    return torchaudio.transforms.Resample(orig, new)


@functools.lru_cache(maxsize=1)
def _processor():
    from transformers import SeamlessM4TFeatureExtractor
    return SeamlessM4TFeatureExtractor()


def split_audio(waveform: torch.Tensor, sample_rate: int, file_name: str, chunk_sec=25, overlap_sec=1,
                target_sr: int = 16000):
    """Resample, split channels, cut overlapped chunks (views) -- audio_processor.py:104-160."""
    if sample_rate != target_sr:
        waveform = _resampler(sample_rate, target_sr)(waveform)
        sample_rate = target_sr
    chunk = int(chunk_sec * sample_rate)
    step = chunk - int(overlap_sec * sample_rate)
    total = waveform.shape[1]
    if waveform.shape[0] == 2:
        agent, client = waveform[0].unsqueeze(0), waveform[1].unsqueeze(0)
        mixed = waveform.mean(dim=0, keepdim=True)
    else:
        agent = client = mixed = waveform
    chunks, start, idx = [], 0, 0
    while start < total:
        end = min(start + chunk, total)
        chunks.append({"file_name": f"{file_name}", "stereo_waveform": mixed[:, start:end],
                       "agent_waveform": agent[:, start:end], "client_waveform": client[:, start:end],
                       "chunk_idx": idx, "start_time": start / sample_rate, "end_time": end / sample_rate})
        idx += 1
        start += step
        if end >= total:
            break
    return chunks, agent, client


def featurise_chunks(chunks: Sequence[Dict], batch_size: int = 16):
    """Length-sorted batches of ``batch_size`` through processor(...) x3 streams
    (audio_processor.py:170-181, :222-224, :253-261)."""
    proc = _processor()
    chunks = sorted(chunks, key=lambda c: int(c["stereo_waveform"].shape[1]))
    out = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i in range(0, len(chunks), batch_size):
            batch = chunks[i:i + batch_size]
            feats = {}
            for key, fld in (("stereo", "stereo_waveform"), ("agent", "agent_waveform"), ("client", "client_waveform")):
                arrays = [c[fld].squeeze().numpy() for c in batch]
                r = proc(arrays, sampling_rate=16000, return_tensors="pt", padding=True)
                feats[key] = (r["input_features"], r["attention_mask"])
            out.append((batch, feats))
    return out


def acoustic_features(waveform: torch.Tensor, sample_rate: int):
    """The 38 scalars of sentiment_analysis.py:807-894 with the same torch / torchaudio calls."""
    import torchaudio
    y = waveform.squeeze(0)
    fl, hop = int(0.025 * sample_rate), int(0.010 * sample_rate)
    if y.numel() < fl:
        return None
    fr = y.unfold(0, fl, hop)
    rms = fr.pow(2).mean(dim=1).sqrt()
    sg = torch.sign(fr)
    zc = ((sg[:, :-1] * sg[:, 1:]) < 0).sum(dim=1).float() / fl
    n_fft = 512
    spec = torch.stft(y, n_fft=n_fft, hop_length=hop, win_length=fl, window=torch.hann_window(fl),
                      return_complex=True)
    mag = spec.abs()
    freqs = torch.linspace(0, sample_rate / 2, mag.shape[0])
    energy = mag.sum(dim=0) + 1e-8
    cen = (mag * freqs[:, None]).sum(dim=0) / energy
    bw = (((freqs[:, None] - cen[None, :]) ** 2 * mag).sum(dim=0) / energy).sqrt()
    cum = mag.cumsum(dim=0)
    thr = 0.85 * (cum[-1] + 1e-8)
    roll = freqs[(cum >= thr).float().argmax(dim=0).long()]
    mfcc = torchaudio.transforms.MFCC(sample_rate=sample_rate, n_mfcc=13,
                                      melkwargs={"n_fft": n_fft, "hop_length": hop, "win_length": fl})(y[None])[0]
    flux = torch.relu(mag[:, 1:] - mag[:, :-1]).sum(dim=0)[None, None]
    ac = torch.nn.functional.conv1d(flux, flux, padding=flux.shape[-1] - 1).squeeze()
    ac[0] = 0
    lag = int(ac.argmax())
    period = lag * hop / sample_rate if lag > 0 else 0.0
    vals = [rms.mean().item(), rms.std().item(), (rms.max() - rms.min()).item(), zc.mean().item(), zc.std().item(),
            cen.mean().item(), cen.std().item(), bw.mean().item(), bw.std().item(),
            roll.mean().item(), roll.std().item(), 60.0 / period if period > 0 else 0.0]
    mm, ms = mfcc.mean(dim=1), mfcc.std(dim=1)
    for i in range(13):
        vals += [mm[i].item(), ms[i].item()]
    return OrderedDict(zip(FEATURE_KEYS, vals))


def front_end_call(call: torch.Tensor, sample_rate: int, acoustic: bool = True, batch_size: int = 16):
    """Everything the hot path does for ONE call on the CPU; returns (n_chunks, call_seconds)."""
    chunks, _, _ = split_audio(call, sample_rate, "call.wav")
    featurise_chunks(chunks, batch_size)
    if acoustic:
        for c in chunks:
            acoustic_features(c["agent_waveform"], 16000)
            acoustic_features(c["client_waveform"], 16000)
    return len(chunks), call.shape[1] / float(sample_rate)


def front_end_group(calls: Sequence[torch.Tensor], sample_rate: int, acoustic: bool = True, batch_size: int = 16):
    """The hot path over a GROUP of calls the way process_file_batch runs it: every call resampled / split, the
    chunks of ALL calls pooled, length-sorted and featurised in batches of ``batch_size``
    (audio_processor.py:170-181), acoustic features per agent / client chunk.  Returns the call seconds."""
    pooled, seconds = [], 0.0
    for i, call in enumerate(calls):
        chunks, _, _ = split_audio(call, sample_rate, f"call{i}.wav")
        pooled += chunks
        seconds += call.shape[1] / float(sample_rate)
    featurise_chunks(pooled, batch_size)
    if acoustic:
        for c in pooled:
            acoustic_features(c["agent_waveform"], 16000)
            acoustic_features(c["client_waveform"], 16000)
    return seconds


def _worker(args):
    """(call seconds processed, seconds spent processing): synthesis of the calls is outside the timed span."""
    seeds, seconds, sample_rate, threads, grouped = args
    torch.set_num_threads(threads)
    from audio_processor_b200 import synth
    total = 0.0
    calls = [synth.make_call(seconds, sample_rate, s, 2) for s in seeds]
    t0 = time.perf_counter()
    if grouped:
        total = front_end_group(calls, sample_rate)
    else:
        for call in calls:
            total += front_end_call(call, sample_rate)[1]
    return total, time.perf_counter() - t0


def timed_sample(n_workers: int, calls_per_worker: int = 1, seconds: float = 60.0, sample_rate: int = 8000,
                 grouped: bool = False, pool=None) -> Tuple[float, float, str]:
    """``n_workers`` host processes x 1 thread (the fbank loop of the reference is single-threaded Python, so
    process parallelism is the fair throughput).  ``grouped``: each worker runs its calls as ONE file batch --
    chunks of all its calls bucketed together (cfg2 shape when calls_per_worker >= 16 and seconds = 300).
    The time is taken INSIDE the workers (the slowest worker's processing span; synthesising the calls is not
    timed).  Returns (audio_hours_per_sec, seconds, description).  ``pool``: a Pool from ``worker_pool``."""
    jobs = [([9000 + w * calls_per_worker + i for i in range(calls_per_worker)], seconds, sample_rate, 1, grouped)
            for w in range(n_workers)]
    own = pool is None
    if own:
        pool = worker_pool(n_workers)
    try:
        res = pool.map(_worker, jobs)
    finally:
        if own:
            pool.terminate()
    audio_s = sum(r[0] for r in res)
    span = max(r[1] for r in res)
    how = "chunks of each process's calls pooled and bucketed in batches of 16" if grouped else "call by call"
    desc = (f"{n_workers * calls_per_worker} synthetic {seconds:.0f} s stereo {sample_rate} Hz calls, "
            f"{n_workers} processes x 1 thread x {calls_per_worker} calls ({how}), full front end "
            f"(resample+split+chunk+fbank x3+acoustic x2), timed inside the workers")
    return audio_s / 3600.0 / span, span, desc


def worker_pool(n_workers: int, sample_rate: int = 8000):
    """A warmed spawn pool: one thread per worker for EVERY library -- torch.set_num_threads does not reach numpy's
    BLAS (the mel matmul of transformers' spectrogram), and n_workers x n_cores BLAS threads oversubscribe the box
    (measured 6.6x slower).  The spawned children inherit the variables."""
    import multiprocessing as mp
    keys = ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")
    saved = {k: os.environ.get(k) for k in keys}
    for k in keys:
        os.environ[k] = "1"
    try:
        pool = mp.get_context("spawn").Pool(n_workers)
        pool.map(_worker, [([1], 2.0, sample_rate, 1, False)] * n_workers)          # warm imports / caches
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return pool
