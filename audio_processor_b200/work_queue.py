"""Multi-GPU sharding of calls: every call is independent and nothing
reduces across calls, so there is no collective on the data path.

* ``shard_indices`` -- the static ``i % n`` deal of the reference's
  ``MultiGPUManager.distribute_batches`` (services/multi_gpu_manager.py:53-65) for
  one-process-per-GPU launches (``torchrun``): rank r takes calls r, r+n, ...
* ``gather_throughput`` -- the only cross-rank traffic: a max over ranks of the step time
  and a sum of the audio seconds, through ``torch.distributed`` (NCCL on GPUs, gloo on CPU).
* ``MultiGpuQueue`` -- in-process alternative: one worker thread per visible GPU pulling
  calls dynamically from a shared queue (no 4-GPU cap, stragglers balance).
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

