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
