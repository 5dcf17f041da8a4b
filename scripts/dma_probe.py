#!/usr/bin/env python
"""Host <-> device copy bandwidth of the box with 1, 2, 4, 8 GPUs copying at once — the ceiling of
bench.py's `e2e` leg, which moves 3.0 GB in and 3.87 GB out per GPU and step through pinned memory.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/dma_probe.py

Every phase k lets the first k ranks run H2D and D2H concurrently on two streams (no kernels);
rank 0 prints one JSON line per phase: aggregate GB/s = k x bytes / max over ranks of the wall time.
"""
import json
import os
import time

import torch
import torch.distributed as dist

H2D_BYTES = 3_000_000_000
D2H_BYTES = 3_867_837_484
REPS = 3


def main():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hin = torch.empty(H2D_BYTES, dtype=torch.uint8).pin_memory()
    hout = torch.empty(D2H_BYTES, dtype=torch.uint8).pin_memory()
    din = torch.empty(H2D_BYTES, dtype=torch.uint8, device="cuda")
    dout = torch.zeros(D2H_BYTES, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    k = 1
    while k <= world:
        for what in ("both", "h2d", "d2h"):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if rank < k:
                for _ in range(REPS):
                    if what != "d2h":
                        with torch.cuda.stream(s1):
                            din.copy_(hin, non_blocking=True)
                    if what != "h2d":
                        with torch.cuda.stream(s2):
                            hout.copy_(dout, non_blocking=True)
                torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            if rank == 0:
                nbytes = (H2D_BYTES if what != "d2h" else 0) + (D2H_BYTES if what != "h2d" else 0)
                sec = dt.item() / REPS
                print(json.dumps({"gpus_copying": k, "direction": what, "s_per_step": sec,
                                  "aggregate_gb_s": k * nbytes / sec / 1e9,
                                  "per_gpu_gb_s": nbytes / sec / 1e9}), flush=True)
        k *= 2
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
