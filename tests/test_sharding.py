"""CPU: record sharding and the optional output gather, world_size 2 over gloo."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200spec import sharding


def test_bounds_partition_everything():
    for n in (0, 1, 7, 1024, 106800):
        for world in (1, 2, 3, 4, 8):
            cuts = [sharding.shard_bounds(n, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        sharding.shard_bounds(10, 2, 2)


def _worker(rank, world, port, n_records):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(n_records * 6, dtype=torch.float32).reshape(n_records, 2, 3)
        lo, hi = sharding.shard_bounds(n_records, world, rank)
        local = full[lo:hi] * 2.0                     # stand-in for the per-rank transform
        got = sharding.gather_spectrograms(local, n_records)
        assert torch.equal(got, full * 2.0)
        assert len(sharding.shard_slice(list(range(n_records)), world, rank)) == hi - lo
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_records", [5, 8])
def test_gather_restores_record_order_world2(n_records):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, n_records), nprocs=2, join=True)


def test_bind_host_to_gpu_is_optional():
    """No NVML / no GPU here: the binding helper reports nothing bound and leaves the affinity alone."""
    import os

    from b200spec.sharding import bind_host_to_gpu

    before = os.sched_getaffinity(0)
    cores = bind_host_to_gpu(0)
    assert isinstance(cores, list)
    assert os.sched_getaffinity(0) == (set(cores) if cores else before)
    os.sched_setaffinity(0, before)
