"""work_queue.MultiGpuProcessQueue on the CPU: two worker processes (no GPU binding), dynamic pull, input order,
per-worker state, error reporting, `.done` / `.failed` markers.  The GPU path is measured by
tools/cfg4_queue_bench.py (profiles/r02_cfg4_queue.md)."""
import os
import time

import pytest

from audio_processor_b200.work_queue import MultiGpuProcessQueue, ProcessedMarkers


def _init(dev, tag):
    return {"dev": dev, "tag": tag, "pid": os.getpid(), "seen": 0}


def _work(state, item):
    state["seen"] += 1
    if item == "boom.wav":
        raise ValueError("cannot decode")
    if isinstance(item, str):
        return item.upper()
    time.sleep(0.02 * (item % 3))                 # uneven items: the queue balances them
    return (item * item, state["tag"], state["pid"], state["seen"])


def test_process_queue_orders_balances_and_reports():
    with MultiGpuProcessQueue([None, None], _init, ("w",)) as q:
        out = q.map(_work, list(range(12)))
        assert [o[0] for o in out] == [i * i for i in range(12)] and {o[1] for o in out} == {"w"}
        pids = {o[2] for o in out}
        assert len(pids) == 2 and os.getpid() not in pids                  # two worker processes, both pulled work
        out2 = q.map(_work, [3, 4])                                         # the same workers (state persists)
        assert {o[2] for o in out2} <= pids and max(o[3] for o in out2) > 1
        with pytest.raises(RuntimeError, match="cannot decode"):
            q.map(_work, ["a.wav", "boom.wav", "c.wav"])
        assert q.map(_work, ["d.wav"]) == ["D.WAV"]                         # the workers survive a failing item


def test_process_queue_markers(tmp_path):
    markers = ProcessedMarkers(tmp_path)
    files = ["a.wav", "boom.wav", "c.wav"]
    with MultiGpuProcessQueue([None, None], _init, ("m",)) as q:
        assert q.map(_work, files, markers=markers) == ["A.WAV", None, "C.WAV"]
        assert (tmp_path / "a.done").exists() and (tmp_path / "c.done").exists()
        assert "cannot decode" in (tmp_path / "boom.failed").read_text()
        assert q.map(_work, files, markers=markers) == [None, None, None]      # .done files are skipped, boom fails again


def _die(state, item):
    if item == 2:
        os._exit(7)                                # a worker that vanishes without a message (segfault, OOM kill)
    time.sleep(0.05)
    return item


def test_process_queue_reports_a_dead_worker_instead_of_hanging():
    q = MultiGpuProcessQueue([None, None], _init, ("d",))
    with pytest.raises(RuntimeError, match="died"):
        q.map(_die, list(range(6)))
    q.close()
