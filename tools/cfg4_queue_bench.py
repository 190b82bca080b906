"""cfg4 as SURVEY.md 8d words it: synthetic 5-minute stereo 8 kHz calls, a pool of distinct calls per GPU in pinned
host memory, file batches pulled through the HOST WORK QUEUE (audio_processor_b200.work_queue.MultiGpuQueue -- the
role of the reference's MultiGPUManager, services/multi_gpu_manager.py:53-65, 103-117) by one worker thread per GPU
in ONE process; run at 1 / 2 / 4 / 8 GPUs, streamed from pinned memory and, separately, with the pool resident in HBM.

    python tools/cfg4_queue_bench.py --gpus 1,2,4,8 [--pool 256] [--batch 64] [--batches-per-gpu 6]

Prints one JSON line per GPU count and variant.  The time is the wall clock around `MultiGpuQueue.map` with every
device synchronised on both sides (several devices: there is no single CUDA event timeline), and beside it the
longest per-device CUDA-event span.  A queue item is one file batch of `--batch` calls (the reference deals batches
of files to GPUs); item i takes calls (i * batch + k) % pool of the pool of the GPU that pulled it.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from audio_processor_b200 import synth  # noqa: E402
from audio_processor_b200.frontend import CallFrontend  # noqa: E402
from audio_processor_b200.work_queue import MultiGpuQueue  # noqa: E402


def make_pool(n: int, seconds: float, seed0: int):
    """n distinct pinned calls: 8 synthesised ones (synth.make_call, 0.4 s each on one core) and circular shifts of
    them by distinct offsets with the channels swapped on odd rounds -- same statistics, different samples."""
    base = [synth.make_call(seconds, 8000, seed0 + i, 2) for i in range(min(8, n))]
    pool = []
    for i in range(n):
        b = base[i % len(base)]
        r = i // len(base)
        x = b if r == 0 else torch.roll(b.flip(0) if r % 2 else b, shifts=7919 * r + 13 * (i % len(base)), dims=1)
        pool.append(x.contiguous().pin_memory())
    return pool


def run(n_gpus: int, pools, dev_pools, a, resident: bool):
    devices = list(range(n_gpus))
    fronts = {d: CallFrontend(torch.device("cuda", d)) for d in devices}
    n_items = a.batches_per_gpu * n_gpus
    items = list(range(n_items))
    spans = {}
    lock = threading.Lock()

    def fn(dev, item):
        src = dev_pools[dev] if resident else pools[dev]
        calls = [src[(item * a.batch + k) % len(src)] for k in range(a.batch)]
        cf = fronts[dev]
        cf.clear_plans()
        with lock:
            first = dev not in spans
            if first:
                e0 = torch.cuda.Event(enable_timing=True); e0.record()
                spans[dev] = [e0, None]
        r = cf.process_calls(calls, 8000, acoustic_to_host=True)
        check = float(r.acoustic_rows[:, 0].sum())          # the batch's host-side result is consumed (synchronises)
        e1 = torch.cuda.Event(enable_timing=True); e1.record()
        spans[dev][1] = e1
        return r.call_seconds, check

    q = MultiGpuQueue(devices)
    q.map(lambda dev, it: fn(dev, it), list(range(n_gpus * 2)))          # warm-up: two batches per GPU
    spans.clear()
    for d in devices:
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    out = q.map(fn, items)
    for d in devices:
        torch.cuda.synchronize(d)
    wall = time.perf_counter() - t0
    audio_s = sum(o[0] for o in out)
    ev_ms = max(e0.elapsed_time(e1) for e0, e1 in spans.values())
    per_dev = {}
    return {"metric": "audio_hours_per_sec", "unit": "audio-hours/s", "workload": "cfg4 through MultiGpuQueue, one process, one worker thread per GPU",
            "input": "resident in HBM" if resident else "pinned host memory (H2D inside the timed region)",
            "n_gpus": n_gpus, "value": audio_s / 3600.0 / wall, "wall_s": wall, "max_device_span_ms": ev_ms,
            "file_batches": n_items, "calls_per_batch": a.batch, "pool_calls_per_gpu": len(pools[0]),
            "call_seconds": a.call_seconds, "audio_hours": audio_s / 3600.0,
            "h2d_bytes": 0 if resident else int(n_items * a.batch * 2 * a.call_seconds * 8000 * 4), **per_dev}


# ---- one worker PROCESS per GPU (work_queue.MultiGpuProcessQueue): what the reference's MultiGPUManager runs ----------
def proc_init(dev, pool_n, seconds, resident_too):
    torch.cuda.set_device(0)                                   # CUDA_VISIBLE_DEVICES = dev was set by the queue
    pool = make_pool(pool_n, seconds, 100 + 1000 * dev)
    state = {"cf": CallFrontend(torch.device("cuda", 0)), "pinned": pool}
    if resident_too:
        state["resident"] = [c.to(torch.device("cuda", 0)) for c in pool]
    return state


def proc_work(state, item):
    idx, batch, resident = item
    src = state["resident"] if resident else state["pinned"]
    calls = [src[(idx * batch + k) % len(src)] for k in range(batch)]
    cf = state["cf"]
    cf.clear_plans()
    r = cf.process_calls(calls, 8000, acoustic_to_host=True)
    return r.call_seconds, float(r.acoustic_rows[:, 0].sum())      # consuming the host-side result synchronises


def run_processes(counts, a):
    from audio_processor_b200.work_queue import MultiGpuProcessQueue
    base = {}
    for n in counts:
        t0 = time.perf_counter()
        with MultiGpuProcessQueue(list(range(n)), proc_init, (a.pool, a.call_seconds, not a.no_resident)) as q:
            q.start(proc_work)
            setup = time.perf_counter() - t0
            for resident in ([False] if a.no_resident else [False, True]):
                q.map(proc_work, [(i, a.batch, resident) for i in range(2 * n)])             # warm-up
                items = [(i, a.batch, resident) for i in range(a.batches_per_gpu * n)]
                t1 = time.perf_counter()
                out = q.map(proc_work, items)
                wall = time.perf_counter() - t1
                audio_s = sum(o[0] for o in out)
                key = "resident in HBM" if resident else "pinned host memory (H2D inside the timed region)"
                line = {"metric": "audio_hours_per_sec", "unit": "audio-hours/s",
                        "workload": "cfg4 through MultiGpuProcessQueue, one worker process per GPU, dynamic pull",
                        "input": key, "n_gpus": n, "value": audio_s / 3600.0 / wall, "wall_s": wall, "file_batches": len(items),
                        "calls_per_batch": a.batch, "pool_calls_per_gpu": a.pool, "call_seconds": a.call_seconds,
                        "audio_hours": audio_s / 3600.0, "worker_setup_s": round(setup, 1),
                        "h2d_bytes": 0 if resident else int(len(items) * a.batch * 2 * a.call_seconds * 8000 * 4)}
                base.setdefault(key, line["value"] / n)
                line["per_gpu_vs_first_row"] = round(line["value"] / n / base[key], 3)
                print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="thread", choices=["thread", "process"])
    ap.add_argument("--gpus", default="1")
    ap.add_argument("--pool", type=int, default=256)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--batches-per-gpu", type=int, default=6)
    ap.add_argument("--call-seconds", type=float, default=300.0)
    ap.add_argument("--no-resident", action="store_true")
    a = ap.parse_args()
    counts = [int(x) for x in a.gpus.split(",")]
    have = torch.cuda.device_count()
    counts = [n for n in counts if n <= have]
    nmax = max(counts)
    if a.mode == "process":
        print(json.dumps({"note": "host", "cpu_count": os.cpu_count(), "gpus_visible": have}), flush=True)
        run_processes(counts, a)
        return
    print(json.dumps({"note": "host", "cpu_count": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)), "gpus_visible": have}), flush=True)
    t0 = time.perf_counter()
    pools = [make_pool(a.pool, a.call_seconds, 100 + 1000 * d) for d in range(nmax)]
    print(json.dumps({"note": "pools", "seconds": round(time.perf_counter() - t0, 1), "pinned_GB": round(nmax * a.pool * 2 * a.call_seconds * 8000 * 4 / 1e9, 1)}), flush=True)
    dev_pools = None
    if not a.no_resident:
        dev_pools = [[c.to(torch.device("cuda", d)) for c in pools[d]] for d in range(nmax)]
    base = {}
    for resident in ([False] if a.no_resident else [False, True]):
        for n in counts:
            line = run(n, pools, dev_pools, a, resident)
            key = line["input"]
            base.setdefault(key, line["value"] / n if n == counts[0] else None)
            if base[key]:
                line["per_gpu_vs_first_row"] = round(line["value"] / n / base[key], 3)
            print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
