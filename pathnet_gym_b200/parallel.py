"""Sharding of the path population over the GPUs of one node (one process per GPU).

The path shards by paths: rank r owns a contiguous slice of the genotypes and their
transitions; parameters and RMSProp slots are replicated; the only data-path exchange is the sum of
the gradient slab.  These helpers are backend-agnostic (NCCL on the GPUs, gloo in the CPU tests)."""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(P_total, rank, world_size):
    """[lo, hi) of the paths owned by `rank`; slices differ by at most one path."""
    base, rem = divmod(int(P_total), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_sum_(t: torch.Tensor):
    """In-place sum over ranks (the shared-module gradient slab)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def allgather_concat(t: torch.Tensor, sizes):
    """Concatenate per-rank slices (scores / genotypes, a few KB) in rank order."""
    rank, ws = world()
    if ws == 1:
        return t
    n = max(sizes)
    pad = torch.zeros((n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    out = [torch.empty_like(pad) for _ in range(ws)]
    dist.all_gather(out, pad)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def bind_to_gpu_numa_node(device_index):
    """Pin this process (and therefore the pinned host buffers it allocates afterwards) to the CPUs NVML reports as
    local to the GPU.  With one process per GPU on a multi-socket host this keeps every rank's staging buffers on the
    memory controller next to its PCIe root; without it the ranks' host-to-device copies share whichever node the
    processes happened to start on.  Returns the CPU list, or None if NVML or the affinity call is unavailable."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = device_index
        if vis:
            parts = vis.split(",")
            if device_index < len(parts) and parts[device_index].isdigit():
                idx = int(parts[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [i for i in range(ncpu) if (int(words[i // 64]) >> (i % 64)) & 1]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:  # noqa: BLE001  (no NVML, restricted container, ...)
        return None
