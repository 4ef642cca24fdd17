"""Concurrent pinned H2D + D2H bandwidth with one process per GPU (torchrun): the ceiling of bench.py's `e2e` leg.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_ranks.py out.json

Every rank copies the bytes of one `e2e` step (655 MB in, 537 MB out) in both directions at once on two streams,
all ranks between barriers; rank 0 writes per-rank and aggregate GB/s."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
bound = []
if world > 1:
    from b200spec.sharding import bind_host_to_gpu
    bound = bind_host_to_gpu(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_in, n_out = 655_360_000, 536_870_912
h_in = torch.empty(n_in, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n_out, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(h2d, d2h, reps=6):
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    mine = time.perf_counter() - t0
    barrier()
    t = torch.tensor([mine], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    nbytes = (n_in if h2d else 0) + (n_out if d2h else 0)
    return nbytes * reps / float(t.item()) / 1e9, nbytes * reps / mine / 1e9


for _ in range(2):
    run(True, True, 2)
res = {"ranks": world, "host_cores": len(os.sched_getaffinity(0)), "cores_bound_rank0": len(bound),
       "bytes_per_step": {"h2d": n_in, "d2h": n_out}}
for name, (a, b) in (("h2d_only", (True, False)), ("d2h_only", (False, True)), ("both", (True, True))):
    slowest, mine = run(a, b)
    res[name + "_GBps_per_gpu_slowest_rank"] = slowest
    res[name + "_GBps_aggregate"] = slowest * world
if rank == 0:
    with open(sys.argv[1] if len(sys.argv) > 1 else "/dev/stdout", "w") as fh:
        json.dump(res, fh, indent=1)
if world > 1:
    dist.destroy_process_group()
