#!/usr/bin/env python
"""Headline benchmark: audio-hours/sec through the waveform front end.

    python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path

One STEP = the whole front end over one batch of synthetic calls per GPU -- BASELINE.json
configs[1]: 64 stereo 8 kHz 5-minute calls -> 16 kHz polyphase resample + agent/client/mixed
split + 25 s chunks with 1 s overlap + fbank for the three streams in length-sorted batches
of 16 with dynamic padding + the 38 acoustic features of every agent and client chunk.
Multi-GPU: calls shard by file across ranks, no collective on the data path (weak scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "audio_hours_per_sec"
UNIT = "audio-hours/s"
FBANK_BYTES_PER_FRAME = 962.0          # 640 B new samples + 320 B features + 2 B mask


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="afe", choices=["afe", "reference"])
    ap.add_argument("--calls", type=int, default=64, help="calls per GPU per step")
    ap.add_argument("--call-seconds", type=float, default=300.0)
    ap.add_argument("--distinct", type=int, default=0, help="distinct synthetic calls per GPU (0: as many as --calls, no reuse)")
    ap.add_argument("--precision", default="f64", choices=["f64", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sub-batch", type=int, default=0, help="calls per pipelined sub-batch (0: library default)")
    return ap.parse_args()


def workload_name(a):
    return (f"cfg2: {a.calls} stereo 8 kHz {a.call_seconds:.0f} s calls per GPU per step -> resample 16 kHz + split + "
            f"25 s/1 s chunks + fbank x3 streams (batches of 16, dynamic padding) + 38 acoustic features x2 speakers")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region: the process is
    started early (it takes a while to print its first line); mark() / stop() bracket the region
    and only lines that arrived inside it are used (the nearest one if the region was short)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines, self.t0 = index, None, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def mark(self):
        self.t0 = time.perf_counter()

    def stop(self):
        t1 = time.perf_counter()
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)                      # let the sample that covers the end of the region arrive
        self.proc.terminate()
        t0 = self.t0 if self.t0 is not None else 0.0
        inside = [ln for ts, ln in self.lines if t0 <= ts <= t1 + 0.1]
        if not inside and self.lines:
            inside = [min(self.lines, key=lambda p: abs(p[0] - t1))[1]]
        sm, mx, reasons = [], [], set()
        for ln in inside:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


CPU_GROUP_CALLS = 16          # calls a CPU worker buckets together: the reference's length sort needs >= batch_size calls


def cpu_sample_config(a):
    """The bounded CPU sample of the cfg2 workload: every worker process runs one file batch of 16 calls of
    --call-seconds, chunks pooled and bucketed in batches of 16 like audio_processor.py:170-181."""
    cores = os.cpu_count() or 1
    return max(1, min(cores, 64)), CPU_GROUP_CALLS, a.call_seconds


def run_reference(a):
    """The reference's CPU implementation of the path (oracle/libref: same torchaudio / transformers / torch calls)
    on all host cores, on the afe arm's workload shape; each step is a bounded sample of it."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    from oracle import libref
    workers, group, seconds = cpu_sample_config(a)
    pool = libref.worker_pool(workers)
    vals, spans, desc = [], [], ""
    try:
        for _ in range(a.warmup):                  # warm-up steps: the same path on one short call per worker
            libref.timed_sample(workers, calls_per_worker=1, seconds=20.0, grouped=True, pool=pool)
        for _ in range(a.steps):
            v, span, desc = libref.timed_sample(workers, calls_per_worker=group, seconds=seconds, grouped=True, pool=pool)
            vals.append(v); spans.append(span)
            if sum(spans) > 100:                   # keep the whole run within a few minutes (a step is ~45 s on 16 vCPUs)
                break
    finally:
        pool.terminate()
    value = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
            "steps": len(vals), "warmup": a.warmup, "ms_per_step": 1e3 * sum(spans) / len(spans),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "calls_per_gpu": a.calls, "call_seconds": a.call_seconds,
                       "sample": desc, "steps_note": f"{len(vals)} of the requested {a.steps} steps: the run stops once ~100 s of CPU work are done"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def fp64_roof(frames_per_launch, launch_ms, num_sms, sm_mhz):
    if not launch_ms or not sm_mhz or a_precision_is_f32():
        return None
    inst_per_frame, issue_rate = 283.0, 0.49
    floor = num_sms * 4 * issue_rate * sm_mhz * 1e6 / inst_per_frame          # frames/s at 100 % FP64 issue
    got = frames_per_launch / (launch_ms * 1e-3)
    return {"bound": "fp64 issue", "fp64_warp_inst_per_frame": inst_per_frame, "issue_per_clk_per_scheduler": issue_rate,
            "floor_frames_per_s": floor, "achieved_frames_per_s": got, "frac": got / floor}


_PRECISION = ["f64"]


def a_precision_is_f32():
    return _PRECISION[0] == "f32"


def ncu_cfg2_summary():
    """Key figures of ONE fbank launch of this very step (cfg2), from the committed `ncu --set full` capture
    (tools/gpu_ncu_bench.sh -> profiles/r02_fbank_cfg2_ncu_summary.json): DRAM bytes and shared-memory wavefronts."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_fbank_cfg2_ncu_summary.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def smem_roof(frames_per_launch, launch_ms, num_sms, sm_mhz, wavefronts_per_frame):
    """The kernel's time against its shared-memory floor: wavefronts per frame (ncu, l1tex__data_pipe_lsu_wavefronts_
    mem_shared) at one wavefront per clock per SM."""
    if not launch_ms or not sm_mhz or not wavefronts_per_frame:
        return None
    floor = num_sms * sm_mhz * 1e6 / wavefronts_per_frame
    got = frames_per_launch / (launch_ms * 1e-3)
    return {"bound": "shared-memory wavefronts", "wavefronts_per_frame": wavefronts_per_frame, "wavefronts_per_clk_per_sm": 1.0,
            "floor_frames_per_s": floor, "achieved_frames_per_s": got, "frac": got / floor}


def run_afe(a):
    import torch
    import torch.distributed as dist
    from audio_processor_b200 import _native, synth
    from audio_processor_b200.frontend import CallFrontend

    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # stdout carries exactly ONE line, the JSON result: anything a library writes to fd 1 on the way (NCCL prints
    # "NCCL version ..." there when NCCL_DEBUG is set) is sent to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # synthetic calls: `distinct` seeded calls per rank, cycled; host copies pinned
    m = max(1, min(a.distinct if a.distinct > 0 else a.calls, a.calls))
    pool = [synth.make_call(a.call_seconds, 8000, 100 + rank * 1000 + i, 2).pin_memory() for i in range(m)]
    host_calls = [pool[i % m] for i in range(a.calls)]
    dev_pool = [c.to(dev) for c in pool]
    dev_calls = [dev_pool[i % m] for i in range(a.calls)]
    prec = _native.PRECISION_F64 if a.precision == "f64" else _native.PRECISION_F32
    _PRECISION[0] = a.precision
    cf = CallFrontend(dev, precision=prec)
    fe = cf.fe
    audio_hours_per_step = a.calls * a.call_seconds / 3600.0
    sb = a.sub_batch if a.sub_batch > 0 else None

    sampler = ClockSampler(local_rank)
    sampler.start()                           # early: nvidia-smi needs a moment before its first line
    for _ in range(max(a.warmup, 3)):
        r = cf.process_calls(dev_calls, 8000, acoustic_to_host=False)
    n_chunks = len(r.chunks)
    frames_per_stream = sum(1 + (e - s - 400) // 160 for _, _, s, e in r.chunks if e - s >= 400)
    stft_frames = sum(1 + (e - s) // 160 for _, _, s, e in r.chunks if e - s >= 400)      # centred STFT of the acoustic path
    del r
    torch.cuda.synchronize(dev)

    # ---- timed region A: inputs resident in HBM ------------------------------------------
    barrier()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        r = cf.process_calls(dev_calls, 8000, acoustic_to_host=False)
        del r
    e1.record()
    barrier()
    ms_dev = e0.elapsed_time(e1)
    # per-kernel CUDA-event times (roofline, kernel_ms_per_step): the same K steps once more with the acoustic chain on
    # the SAME stream as fbank -- in the timed steps above it runs on a second stream, where a kernel's event pair also
    # spans its wait for free SMs
    cf.kTwoStreams = False
    fe.profile(True)
    for _ in range(a.steps):
        r = cf.process_calls(dev_calls, 8000, acoustic_to_host=False)
        del r
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    prof = fe.profile_read()
    fe.profile(False)
    cf.kTwoStreams = True

    # ---- timed region B: end to end from pinned host buffers -----------------------------
    for _ in range(2):
        r = cf.process_calls(host_calls, 8000, acoustic_to_host=True, sub_batch=sb)
    del r
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    f0.record()
    checksum = 0.0
    for _ in range(a.steps):
        cf.clear_plans()                                      # the host plans the layout every step, as for ragged batches
        r = cf.process_calls(host_calls, 8000, acoustic_to_host=True, sub_batch=sb)
        checksum += float(r.acoustic_rows[:, 0].sum())        # the step's host-side result is consumed
        del r
    f1.record()
    barrier()
    ms_e2e = max(f0.elapsed_time(f1), 1e3 * (time.perf_counter() - t0))

    # ---- the floor under region B: the same pinned buffers, upload only (one cudaMemcpyAsync per call on one
    # stream, all ranks at once, so GPUs that share a host uplink show it) ---------------------------------
    sink = [torch.empty_like(c, device=dev) for c in pool]
    h2d_ms = []
    for rep_i in range(4):
        barrier()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0.record()
        for i in range(a.calls):
            sink[i % m].copy_(host_calls[i], non_blocking=True)
        h1.record()
        barrier()
        if rep_i:
            h2d_ms.append(h0.elapsed_time(h1))
    ms_h2d = min(h2d_ms)
    del sink

    # ---- timed region C: end to end from pinned 16-bit wav file images ---------
    # each call is the image of a 16-bit stereo wav file in pinned memory (44-byte RIFF header + interleaved
    # frames); every step walks the headers (afe_wav_parse) and hands the frames over where they lie
    from audio_processor_b200 import wav
    wav_pool = [synth.wav_image(c, 8000).pin_memory() for c in pool]
    wav_files = [wav_pool[i % m] for i in range(a.calls)]
    pcm_calls = [wav.load_pcm16(f)[0] for f in wav_files]
    for _ in range(2):
        r = cf.process_calls(pcm_calls, 8000, acoustic_to_host=True, sub_batch=sb)
    del r
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    g0.record()
    checksum_pcm = 0.0
    for _ in range(a.steps):
        cf.clear_plans()
        pcm_calls = [wav.load_pcm16(f)[0] for f in wav_files]
        r = cf.process_calls(pcm_calls, 8000, acoustic_to_host=True, sub_batch=sb)
        checksum_pcm += float(r.acoustic_rows[:, 0].sum())
        del r
    g1.record()
    barrier()
    ms_pcm = max(g0.elapsed_time(g1), 1e3 * (time.perf_counter() - t0))

    times = torch.tensor([ms_dev, ms_e2e, ms_pcm, ms_h2d], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_dev, ms_e2e, ms_pcm, ms_h2d = float(times[0]), float(times[1]), float(times[2]), float(times[3])

    if rank == 0:
        total_hours = audio_hours_per_step * a.steps * world
        value = total_hours / (ms_dev / 1e3)
        e2e_value = total_hours / (ms_e2e / 1e3)
        fb_ms, fb_n = prof["fbank"]
        peak, peak_src = measured_peak()
        # the three streams' frames of a step, spread over however many launches the step used
        frames_per_launch = 3.0 * frames_per_stream * a.steps / max(fb_n, 1)
        achieved = frames_per_launch * FBANK_BYTES_PER_FRAME / (fb_ms / max(fb_n, 1) * 1e-3) / 1e9 if fb_ms > 0 else None
        ncu = ncu_cfg2_summary()
        traffic_pf = ncu.get("dram_bytes_per_frame")
        kernel_ms = {k: round(v[0] / a.steps, 4) for k, v in prof.items() if v[1]}
        # the other kernels against the same HBM roof (algorithmic bytes of SURVEY.md 8d per step and GPU):
        # resample 8 kHz stereo 256 000 B per call-second; acoustic tile 640 B per 10 ms frame of each speaker;
        # MFCC pass: the 128 mel-dB floats of every frame, written by the tile kernel and read back once
        # acoustic: 640 B per 10 ms frame of each speaker + 152 B per waveform -- the 128 mel-dB floats the
        # tile kernel writes and the MFCC kernel reads back (1 024 B per frame) are an intermediate: overhead, not credit
        acoustic_bytes = 2.0 * stft_frames * 640.0 + 2.0 * n_chunks * 152.0
        step_bytes = {"resample": 256000.0 * a.calls * a.call_seconds, "acoustic_tile": acoustic_bytes}
        other = {}
        for k, nbytes in step_bytes.items():
            if k in prof and prof[k][0] > 0:
                gbps = nbytes / (prof[k][0] / a.steps * 1e-3) / 1e9
                other[k] = {"achieved": round(gbps, 1), "frac": round(gbps / peak, 4), "unit": "GB/s",
                            "algorithmic_bytes_per_step": nbytes}
        chain_ms = sum(prof[k][0] for k in ("acoustic_tile", "acoustic_max", "acoustic_mfcc", "acoustic_final") if k in prof) / a.steps
        if chain_ms > 0:
            gbps = acoustic_bytes / (chain_ms * 1e-3) / 1e9
            other["acoustic_chain"] = {"achieved": round(gbps, 1), "frac": round(gbps / peak, 4), "unit": "GB/s",
                                       "algorithmic_bytes_per_step": acoustic_bytes, "ms_per_step": round(chain_ms, 4),
                                       "overhead_bytes_per_step": 2.0 * stft_frames * 1024.0,
                                       "overhead_note": "mel-dB rows written by the tile kernel and read back by the MFCC kernel "
                                                        "(the global top_db maximum forces the second pass)"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if a.precision == "f64" else "f32", "data": "synthetic",
            "config": {"workload": workload_name(a), "calls_per_gpu": a.calls, "call_seconds": a.call_seconds,
                       "chunks_per_step_per_gpu": n_chunks, "distinct_calls": m, "parallelism": f"dp{world} by file",
                       "l2": f"{m} distinct calls per GPU = {m * 2 * a.call_seconds * 8000 * 4 / 1e6:.0f} MB of distinct input per step "
                             f"({'larger than' if m * 2 * a.call_seconds * 8000 * 4 > 126e6 * 2 else 'NOT larger than twice'} the 126 MB L2); "
                             "every kernel's input was written by the previous kernel or step, no explicit flush",
                       "fbank_precision": a.precision, "kernel_ms_per_step": kernel_ms,
                       "kernel_ms_note": "per-kernel CUDA-event times from a second pass of the same steps with the acoustic chain "
                                         "on the fbank stream (in the timed steps it overlaps the fbank kernels' tail on a second stream)"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / a.steps,
                    "h2d_bytes_per_step": int(a.calls * 2 * a.call_seconds * 8000 * 4),
                    "d2h_bytes_per_step": int(2 * n_chunks * (38 * 4 + 4)), "checksum": checksum,
                    "host_planning": "inside the timed steps (plan cache cleared every step)",
                    "h2d_floor_ms": ms_h2d,
                    "h2d_floor_note": "upload only of the same pinned buffers, one cudaMemcpyAsync per call, all ranks at once; "
                                      "best of 3, max over ranks",
                    "h2d_floor_frac": ms_h2d / (ms_e2e / a.steps)},
            "e2e_wav16": {"value": total_hours / (ms_pcm / 1e3), "unit": UNIT, "ms_per_step": ms_pcm / a.steps,
                          "h2d_bytes_per_step": int(a.calls * 2 * a.call_seconds * 8000 * 2),
                          "d2h_bytes_per_step": int(2 * n_chunks * (38 * 4 + 4)), "checksum": checksum_pcm,
                          "note": "same calls as pinned images of 16-bit stereo wav files: header walk on the host, interleaved "
                                  "frames cross PCIe as they lie, de-interleave + x/32768 fused into the resampler"},
            "gpu_launches": int(sum(v[1] for v in prof.values())),
            "roofline": {"bound": "hbm", "kernel": "phased_kernel<FbankKernel<%s>>" % ("double" if a.precision == "f64" else "float"),
                         "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": (achieved / peak) if achieved else None,
                         "traffic": (traffic_pf * frames_per_launch) if traffic_pf else None,
                         "traffic_source": ("profiles/r02_fbank_cfg2_ncu_summary.json: dram__bytes_read+write of one fbank launch of "
                                            "this step under ncu --set full, per frame x frames per launch") if traffic_pf else None,
                         "algorithmic_bytes_per_launch": frames_per_launch * FBANK_BYTES_PER_FRAME,
                         "launches_timed": fb_n, "avg_launch_ms": fb_ms / max(fb_n, 1),
                         "note": "fp64-issue-bound, not HBM-bound: see DESIGN.md section 5"},
            # the pipe that actually bounds the kernel: 283 FP64 warp-instructions per frame at the
            # measured 0.49 warp-instructions/clk/scheduler (profiles/r01_pipe_probe_b200.txt), 4 schedulers x SMs
            "roofline_fp64": fp64_roof(frames_per_launch, fb_ms / max(fb_n, 1), fe.num_sms, clocks.get("sm_mhz")),
            # ... and the one that binds first: shared-memory wavefronts per frame from the same ncu capture
            "roofline_smem": smem_roof(frames_per_launch, fb_ms / max(fb_n, 1), fe.num_sms, clocks.get("sm_mhz"),
                                       ncu.get("smem_wavefronts_per_frame")),
            "roofline_other_kernels": other,
        }
        if world == 1 and not a.no_cpu_baseline:
            from oracle import libref
            workers, group, seconds = cpu_sample_config(a)
            v, span, desc = libref.timed_sample(workers, calls_per_worker=group, seconds=seconds, grouped=True)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": workers, "kind": "port", "sample": desc, "wall_s": span}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    os.close(json_fd)


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_afe(a)


if __name__ == "__main__":
    main()
