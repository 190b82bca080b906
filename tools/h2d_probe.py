"""Raw pinned host->device bandwidth of the bench's e2e inputs (64 calls of 19.2 MB) on this box."""
import os, sys, time
import torch
dev = torch.device("cuda", 0)
pool = [torch.randn(2, 2_400_000).pin_memory() for _ in range(8)]
calls = [pool[i % 8] for i in range(64)]
dst = torch.empty(64, 2, 2_400_000, device=dev)
s = torch.cuda.Stream(dev)
for label, chunks in (("64 x 19.2 MB", calls),):
    for rep in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.cuda.stream(s):
            for i, c in enumerate(chunks):
                dst[i].copy_(c, non_blocking=True)
        s.synchronize()
        dt = time.perf_counter() - t0
        print(f"{label}: {1e3*dt:.2f} ms, {64*19.2e6/dt/1e9:.1f} GB/s", flush=True)
big = torch.randn(64, 2, 2_400_000).pin_memory()
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); dst.copy_(big, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"one 1.23 GB copy: {1e3*dt:.2f} ms, {1.2288e9/dt/1e9:.1f} GB/s", flush=True)
for nstream in (2, 3, 4):
    ss = [torch.cuda.Stream(dev) for _ in range(nstream)]
    for rep in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i, c in enumerate(calls):
            with torch.cuda.stream(ss[i % nstream]):
                dst[i].copy_(c, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"{nstream} streams, 64 x 19.2 MB: {1e3*dt:.2f} ms, {64*19.2e6/dt/1e9:.1f} GB/s", flush=True)
# halves of each call on one stream (128 copies)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(s):
        for i, c in enumerate(calls):
            dst[i, 0].copy_(c[0], non_blocking=True); dst[i, 1].copy_(c[1], non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"1 stream, 128 x 9.6 MB: {1e3*dt:.2f} ms, {64*19.2e6/dt/1e9:.1f} GB/s", flush=True)
