"""Record sharding across GPUs: one process per GPU, contiguous slices, no collective.

Every (record, channel) row of mode R and every record of mode N is independent (the
normalisation of /root/reference/root/src/utils/DiffEEG_utils.py:72-74 is per row), so rank r
of G simply takes records ``[bounds(r), bounds(r+1))``.  The only optional exchange is a final
all-gather of the outputs for a consumer that wants them replicated.
"""
from __future__ import annotations

import torch


def shard_bounds(n_records: int, world_size: int, rank: int) -> tuple[int, int]:
    """Contiguous slice of rank ``rank``; the remainder goes to the low ranks."""
    if world_size <= 0 or not 0 <= rank < world_size or n_records < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(n_records, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_slice(records, world_size: int, rank: int):
    lo, hi = shard_bounds(len(records), world_size, rank)
    return records[lo:hi]


def gather_spectrograms(local: torch.Tensor, n_records: int, group=None) -> torch.Tensor:
    """All-gather per-rank outputs back into record order (NCCL on GPUs, gloo on CPU).

    Shards may differ by one record, so every rank pads to the largest shard, one
    ``all_gather_into_tensor`` moves the data, and the padding is dropped."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = [shard_bounds(n_records, world, r) for r in range(world)]
    biggest = max(hi - lo for lo, hi in sizes)
    padded = local
    if local.shape[0] < biggest:
        pad = torch.zeros((biggest - local.shape[0],) + tuple(local.shape[1:]),
                          dtype=local.dtype, device=local.device)
        padded = torch.cat([local, pad], dim=0)
    out = torch.empty((world * biggest,) + tuple(local.shape[1:]), dtype=local.dtype,
                      device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous(), group=group)
    if all(hi - lo == biggest for lo, hi in sizes):
        return out                      # equal shards: the collective's buffer IS the result, no second copy
    parts =[out[r * biggest: r * biggest + (hi - lo)] for r, (lo, hi) in enumerate(sizes)]
    return torch.cat(parts, dim=0)


def bind_host_to_gpu(device_index: int) -> list[int]:
    """Pin this process to the CPU cores next to GPU ``device_index`` (its NUMA node), so that the
    pinned host buffers it allocates afterwards (first touch) and the copy threads sit on the
    socket the GPU's PCIe link hangs off.  One process per GPU: without it eight ranks share
    whatever node the launcher started them on and the host-buffer path runs at a fraction of the
    link rate.  Returns the cores bound to ([] when NVML or the affinity call is unavailable;
    binding is an optimisation, never a requirement)."""
    import os

    try:
        import pynvml

        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        if visible:                                           # NVML enumerates physical devices
            ids = [v.strip() for v in visible.split(",") if v.strip()]
            tok = ids[device_index]
            handle = pynvml.nvmlDeviceGetHandleByUUID(tok) if tok.startswith("GPU-") else \
                pynvml.nvmlDeviceGetHandleByIndex(int(tok))
        else:
            handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        allowed = os.sched_getaffinity(0)
        cores = [c for c in range(n_cpu) if (words[c // 64] >> (c % 64)) & 1 and c in allowed]
        if not cores:
            return []
        os.sched_setaffinity(0, cores)
        return cores
    except Exception:
        return []
