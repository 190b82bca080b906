"""N > 1 host logic on CPU: world_size-2 gloo ranks shard calls with no data-path
collective and agree on (sum of audio seconds, max step time)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from audio_processor_b200 import work_queue


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _rank_main(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = work_queue.shard_indices(7, rank, world)
    secs = sum(10.0 * (i + 1) for i in mine)               # call i lasts 10*(i+1) s
    total, worst = work_queue.gather_throughput(secs, 100.0 + 50.0 * rank)
    ret[rank] = (mine, total, worst)
    dist.destroy_process_group()


def test_two_gloo_ranks_shard_by_file():
    world, port = 2, _free_port()
    ret = mp.Manager().dict()
    mp.spawn(_rank_main, args=(world, port, ret), nprocs=world, join=True)
    assert ret[0][0] == [0, 2, 4, 6] and ret[1][0] == [1, 3, 5]
    assert ret[0][1] == ret[1][1] == sum(10.0 * (i + 1) for i in range(7))
    assert ret[0][2] == ret[1][2] == 150.0


def test_shard_indices_cover_everything_once():
    for n in (0, 1, 5, 64):
        for w in (1, 2, 4, 8):
            got = sorted(i for r in range(w) for i in work_queue.shard_indices(n, r, w))
            assert got == list(range(n))
    with pytest.raises(ValueError):
        work_queue.shard_indices(4, 2, 2)


def test_multi_gpu_queue_dynamic_pull_without_devices():
    q = work_queue.MultiGpuQueue(devices=[0, 1, 2])
    out = q.map(lambda dev, x: (x * x), list(range(20)), set_device=False)
    assert out == [x * x for x in range(20)]
    with pytest.raises(ZeroDivisionError):
        q.map(lambda dev, x: 1 // x, [1, 0, 2], set_device=False)


def test_processed_markers_and_per_file_retries(tmp_path):
    """The .done / .failed markers of DataProcessor (services/pipeline.py:280-315): names, contents, resume."""
    import json
    from datetime import datetime
    markers = work_queue.ProcessedMarkers(tmp_path / "markers")
    files = [tmp_path / f"call_{i}.wav" for i in range(6)]
    markers.mark_file_processed(files[1])                       # finished by an earlier run
    attempts = {}

    def fn(dev, path):
        attempts[path.stem] = attempts.get(path.stem, 0) + 1
        if path.stem == "call_3":
            raise RuntimeError("decode failed")
        if path.stem == "call_4" and attempts[path.stem] < 2:
            raise IOError("transient")
        return path.stem.upper()
    out = work_queue.MultiGpuQueue(devices=[0, 1]).map(fn, files, set_device=False, markers=markers, max_retries=1)
    assert out == ["CALL_0", None, "CALL_2", None, "CALL_4", "CALL_5"]
    assert "call_1" not in attempts and attempts["call_3"] == 2 and attempts["call_4"] == 2
    done = sorted(p.name for p in (tmp_path / "markers").glob("*.done"))
    assert done == ["call_0.done", "call_1.done", "call_2.done", "call_4.done", "call_5.done"]
    datetime.fromisoformat((tmp_path / "markers" / "call_0.done").read_text())
    failed = json.loads((tmp_path / "markers" / "call_3.failed").read_text())
    assert failed["error"] == "decode failed" and set(failed) == {"timestamp", "error"}
    assert markers.is_already_processed("anywhere/else/call_5.ogg") and not markers.is_already_processed(files[3])
