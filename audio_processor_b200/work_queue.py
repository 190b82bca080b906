"""Multi-GPU sharding of calls: every call is independent and nothing
reduces across calls, so there is no collective on the data path.

* ``shard_indices`` -- the static ``i % n`` deal of the reference's
  ``MultiGPUManager.distribute_batches`` (services/multi_gpu_manager.py:53-65) for
  one-process-per-GPU launches (``torchrun``): rank r takes calls r, r+n, ...
* ``gather_throughput`` -- the only cross-rank traffic: a max over ranks of the step time
  and a sum of the audio seconds, through ``torch.distributed`` (NCCL on GPUs, gloo on CPU).
* ``MultiGpuQueue`` -- in-process alternative: one worker thread per visible GPU pulling
  calls dynamically from a shared queue (no 4-GPU cap, stragglers balance).
* ``MultiGpuProcessQueue`` -- the same dynamic pull with one worker PROCESS per GPU, which is what the
  reference's ``MultiGPUManager`` runs (one process per GPU, ``CUDA_VISIBLE_DEVICES`` set, ``gpu_index`` 0:
  multi_gpu_manager.py:298, 341).  Measured on 8 x B200 (profiles/r02_cfg4_queue.md): the thread variant
  stops scaling at 2 GPUs because the Python side of every batch (planning, ctypes calls) holds the GIL;
  processes scale like ``torchrun``.
* ``ProcessedMarkers`` -- the ``<stem>.done`` / ``<stem>.failed`` marker files of
  ``DataProcessor`` (services/pipeline.py:280-315), same names and contents, so a run can be
  resumed by either implementation.
"""
from __future__ import annotations

import json
import queue
import threading
from datetime import datetime
from pathlib import Path
from typing import Callable, List, Optional, Sequence, Tuple, Union


def shard_indices(n_items: int, rank: int, world_size: int) -> List[int]:
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    return list(range(rank, n_items, world_size))


def gather_throughput(local_audio_seconds: float, local_ms: float, device=None) -> Tuple[float, float]:
    """(total audio seconds over all ranks, max step time in ms over all ranks)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(local_audio_seconds), float(local_ms)
    t_sum = torch.tensor([local_audio_seconds], dtype=torch.float64, device=device)
    t_max = torch.tensor([local_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t_sum, op=dist.ReduceOp.SUM)
    dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
    return float(t_sum[0]), float(t_max[0])


class MultiGpuQueue:
    """``results = MultiGpuQueue(devices).map(fn, items)`` with ``fn(device_index, item)``
    executed on a worker thread bound to that device; results keep the input order."""

    def __init__(self, devices: Optional[Sequence[int]] = None):
        if devices is None:
            import torch
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise RuntimeError("MultiGpuQueue needs at least one device")
        self.devices = list(devices)

    def map(self, fn: Callable, items: Sequence, set_device: bool = True, markers: "Optional[ProcessedMarkers]" = None,
            max_retries: int = 0) -> List:
        """Without ``markers`` the first exception stops that worker and is re-raised after the join.  With
        ``markers`` (items are file paths) the reference's per-file contract applies: files already marked
        ``.done`` are skipped (result ``None``), a file is retried ``max_retries`` times, then marked ``.failed``
        with the error text and the other files carry on; successes are marked ``.done``."""
        q: "queue.Queue[Tuple[int, object]]" = queue.Queue()
        for i, it in enumerate(items):
            if markers is not None and markers.is_already_processed(it):
                continue
            q.put((i, it))
        out: List = [None] * len(items)
        errors: List[BaseException] = []
        if markers is not None:
            inner = fn

            def fn(dev, it):                        # noqa: F811 -- per-file retry + marker wrapper
                err: Optional[BaseException] = None
                for _ in range(max(0, int(max_retries)) + 1):
                    try:
                        res = inner(dev, it)
                        markers.mark_file_processed(it)
                        return res
                    except Exception as e:
                        err = e
                markers.mark_file_failed(it, str(err))
                return None

        def work(dev: int):
            if set_device:
                import torch
                torch.cuda.set_device(dev)
            while True:
                try:
                    i, it = q.get_nowait()
                except queue.Empty:
                    return
                try:
                    out[i] = fn(dev, it)
                except BaseException as e:          # surfaced to the caller after the join
                    errors.append(e)
                    return

        threads = [threading.Thread(target=work, args=(d,), daemon=True) for d in self.devices]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return out


def _process_worker(dev, init_fn, init_args, fn, tasks, results):
    """Body of one MultiGpuProcessQueue worker: binds the process to its GPU the way the reference does
    (CUDA_VISIBLE_DEVICES before the first CUDA call, device index 0 inside), builds its state once, then pulls
    ``(index, item)`` pairs until the ``None`` sentinel."""
    import os
    import traceback
    try:
        if dev is not None:
            os.environ["CUDA_VISIBLE_DEVICES"] = str(dev)
        state = init_fn(dev, *init_args) if init_fn is not None else None
        results.put(("ready", dev, None))
    except BaseException as e:                       # noqa: BLE001
        results.put(("dead", dev, f"{e!r}\n{traceback.format_exc()}"))
        return
    while True:
        task = tasks.get()
        if task is None:
            return
        i, item = task
        try:
            results.put(("ok", i, fn(state, item)))
        except BaseException as e:                   # noqa: BLE001
            results.put(("err", i, f"{e!r}\n{traceback.format_exc()}"))


class MultiGpuProcessQueue:
    """Host work queue over one worker process per GPU, dynamic pull, any GPU count::

        with MultiGpuProcessQueue(devices, init_fn, init_args) as q:     # init_fn(dev, *init_args) -> state, once per worker
            results = q.map(fn, items)                                   # fn(state, item); results keep the input order

    ``init_fn`` / ``fn`` are module-level functions (the workers are spawned).  Inside a worker the GPU is device 0
    (``CUDA_VISIBLE_DEVICES`` = the physical index, as multi_gpu_manager.py:298 does); a ``None`` device binds nothing
    (CPU tests).  An item that raises is reported after the map as ``RuntimeError`` carrying the worker's traceback;
    the other items still run.  With ``markers`` the reference's per-file contract applies in the parent: ``.done``
    files are skipped, successes marked ``.done``, failures ``.failed`` (no exception then)."""

    def __init__(self, devices: Optional[Sequence[Optional[int]]] = None, init_fn: Optional[Callable] = None,
                 init_args: tuple = ()):
        if devices is None:
            import torch
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise RuntimeError("MultiGpuProcessQueue needs at least one device")
        self.devices = list(devices)
        self._init = (init_fn, tuple(init_args))
        self._procs: list = []
        self._tasks = self._results = None
        self._fn = None

    def start(self, fn: Callable) -> "MultiGpuProcessQueue":
        import multiprocessing as mp
        if self._procs:
            if fn is not self._fn:
                raise RuntimeError("the workers were started with another function")
            return self
        ctx = mp.get_context("spawn")
        self._tasks, self._results, self._fn = ctx.Queue(), ctx.Queue(), fn
        for d in self.devices:
            p = ctx.Process(target=_process_worker, args=(d, self._init[0], self._init[1], fn, self._tasks, self._results),
                            daemon=True)
            p.start()
            self._procs.append(p)
        for _ in self.devices:
            kind, dev, msg = self._next_result()
            if kind != "ready":
                self.close()
                raise RuntimeError(f"worker of device {dev} failed to start: {msg}")
        return self

    def map(self, fn: Callable, items: Sequence, markers: "Optional[ProcessedMarkers]" = None) -> List:
        self.start(fn)
        todo = [(i, it) for i, it in enumerate(items) if not (markers is not None and markers.is_already_processed(it))]
        for t in todo:
            self._tasks.put(t)
        out: List = [None] * len(items)
        errors = []
        for _ in todo:
            kind, i, payload = self._next_result()
            if kind == "ok":
                out[i] = payload
                if markers is not None:
                    markers.mark_file_processed(items[i])
            elif markers is not None:
                markers.mark_file_failed(items[i], str(payload).splitlines()[0])
            else:
                errors.append((i, payload))
        if errors:
            raise RuntimeError(f"item {errors[0][0]} failed in its worker: {errors[0][1]}")
        return out

    def _next_result(self):
        """Next message of a worker; a worker that died without one (segfault, OOM kill) must not hang the parent."""
        import queue as _queue
        while True:
            try:
                return self._results.get(timeout=1.0)
            except _queue.Empty:
                dead = [p for p in self._procs if not p.is_alive()]
                if dead:
                    codes = ", ".join(str(p.exitcode) for p in dead)
                    self.close()
                    raise RuntimeError(f"{len(dead)} worker process(es) died (exit code {codes}) with work outstanding")

    def close(self) -> None:
        for _ in self._procs:
            try:
                self._tasks.put(None)
            except Exception:
                pass
        for p in self._procs:
            p.join(timeout=10)
            if p.is_alive():
                p.terminate()
        self._procs = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class ProcessedMarkers:
    """Idempotence markers of ``DataProcessor`` (services/pipeline.py:280-315): ``<stem>.done`` holds an ISO
    timestamp, ``<stem>.failed`` a JSON object ``{"timestamp", "error"}``.  Every method swallows I/O errors
    like the reference (a marker that cannot be written must not fail the file)."""

    def __init__(self, directory: Union[str, Path]):
        self.dir = Path(directory)
        self.dir.mkdir(parents=True, exist_ok=True)

    def is_already_processed(self, file_path: Union[str, Path]) -> bool:
        try:
            return (self.dir / f"{Path(file_path).stem}.done").exists()
        except Exception:
            return False

    def mark_file_processed(self, file_path: Union[str, Path]) -> None:
        try:
            (self.dir / f"{Path(file_path).stem}.done").write_text(datetime.now().isoformat())
        except Exception:
            pass

    def mark_file_failed(self, file_path: Union[str, Path], error: str = "") -> None:
        try:
            content = {"timestamp": datetime.now().isoformat(), "error": error}
            (self.dir / f"{Path(file_path).stem}.failed").write_text(json.dumps(content))
        except Exception:
            pass

