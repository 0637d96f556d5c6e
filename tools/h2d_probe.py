#!/usr/bin/env python
"""Pinned-host -> device copy bandwidth of this box (one process per visible GPU, concurrently)."""
import sys, time, torch, torch.multiprocessing as mp

def work(rank, n, out):
    torch.cuda.set_device(rank)
    x = torch.empty(256, 3, 224, 224).pin_memory()
    d = torch.empty_like(x, device=f"cuda:{rank}")
    for _ in range(3):
        d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(f"procs {n} rank {rank}: {x.numel() * 4 / dt / 1e9:.1f} GB/s ({dt * 1e3:.2f} ms per 154 MB)", flush=True)

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    mp.spawn(work, args=(n, None), nprocs=n, join=True)
