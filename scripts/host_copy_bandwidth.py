"""Aggregate host<->device copy bandwidth of the box when N ranks copy at once (torchrun): the ceiling of any
end-to-end number that moves its inputs and outputs through pinned host memory every step."""
import os, sys, time
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29577")
dist.init_process_group("nccl", device_id=torch.device("cuda", local), rank=rank, world_size=world)
MB = 1 << 20
for nbytes in (16 * MB, 64 * MB):
    h_in, h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory(), torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty(nbytes, dtype=torch.uint8, device="cuda"), torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for mode in ("h2d", "d2h", "both"):
        def go(n):
            for _ in range(n):
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s1):
                        d_in.copy_(h_in, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s2):
                        h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
        go(3)
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter(); go(20); dt = time.perf_counter() - t0
        t = torch.tensor([dt], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per = (2 if mode == "both" else 1) * nbytes * 20 / float(t) / 1e9
        if rank == 0:
            print(f"{world} ranks, {nbytes // MB} MiB per copy, {mode}: {per:.1f} GB/s per rank, {per * world:.1f} GB/s aggregate", flush=True)
dist.destroy_process_group()
