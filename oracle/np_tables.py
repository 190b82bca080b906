"""TEST INFRASTRUCTURE -- integer / float64 contracts of the hot path's bookkeeping.

* ``chunk_table``      services/audio_processor.py:115-158 (25 s chunks, 1 s overlap)
* ``merge_intervals``  services/speech_segment.py:366-383
* ``segment_table``    services/speech_segment.py:385-418 (pad, window, int() truncation)
* ``bucket_order``     services/audio_processor.py:170-181 (stable sort by length, slices)

Everything here must be reproduced BIT-EXACTLY by the product.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def chunk_table(total_samples: int, sample_rate: int, chunk_duration_sec=25, overlap_sec=1):
    """List of ``(start, end, start_time, end_time)``; an empty waveform gives no chunks."""
    chunk = int(chunk_duration_sec * sample_rate)
    overlap = int(overlap_sec * sample_rate)
    step = chunk - overlap
    out = []
    start = 0
    while start < total_samples:
        end = min(start + chunk, total_samples)
        out.append((start, end, start / sample_rate, end / sample_rate))
        start += step
        if end >= total_samples:
            break
    return out


def merge_intervals(timeline: Sequence[Tuple[float, float]], gap_threshold: float):
    if not timeline:
        return []
    merged = []
    cs, ce = timeline[0]
    for s, e in timeline[1:]:
        if s - ce <= gap_threshold:
            ce = e
        else:
            merged.append((cs, ce))
            cs, ce = s, e
    merged.append((cs, ce))
    return merged


def segment_table(num_samples: int, sample_rate: int, segments: Sequence[Tuple[float, float]],
                  padding: float, max_len_sec: float = 20.0, overlap: float = 0.5):
    """List of ``(start_sample, end_sample, start_sec, end_sec)`` in emission order."""
    total = num_samples / sample_rate
    out = []
    for start, end in segments:
        ps = max(0.0, start - padding)
        pe = min(total, end + padding)
        cur = ps
        while cur < pe:
            we = min(pe, cur + max_len_sec)
            out.append((int(cur * sample_rate), int(we * sample_rate), cur, we))
            if we >= pe:
                break
            cur = max(cur + max_len_sec - overlap, cur + 0.1)
    return out


def bucket_order(lengths: Sequence[int], batch_size: int, enable: bool = True) -> List[List[int]]:
    """Index batches after the reference's stable length sort."""
    order = sorted(range(len(lengths)), key=lambda i: int(lengths[i])) if enable else list(range(len(lengths)))
    return [order[i:i + batch_size] for i in range(0, len(order), batch_size)]
