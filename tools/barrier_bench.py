"""Device-side slab barrier vs a one-element NCCL all-reduce, back to back on every rank (torchrun, one box)."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_sys_b200 import parallel as PL  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ex = PL.PeerExchange(rows_cap=1024)
flag = torch.zeros(1, device=dev)
counts = torch.arange(9, device=dev, dtype=torch.int32)
n = 200
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, fn in (("slab barrier", lambda: ex.barrier()), ("slab barrier + counts", lambda: ex.barrier(counts)),
                 ("nccl all_reduce(1)", lambda: dist.all_reduce(flag))):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    th = time.perf_counter() - t0
    torch.cuda.synchronize()
    if rank == 0:
        print(f"{name:24s} world {world}: {e0.elapsed_time(e1) / n * 1e3:7.1f} us device, {th / n * 1e6:7.1f} us host enqueue", flush=True)
ex.close()
dist.destroy_process_group()
