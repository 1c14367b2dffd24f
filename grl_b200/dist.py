"""Host-side multi-GPU plumbing: agents shard trivially across ranks (no data-path collective); the only
exchange is the final gather of per-agent returns and a reduction of counters/timings."""
import torch
import torch.distributed as dist


def shard(rank, world, agents_per_rank):
    """Global agent ids owned by `rank`: [first, first + count). Weak scaling: fixed agents per rank."""
    return rank * agents_per_rank, agents_per_rank


def finalize(timings, counters, per_agent, device):
    """MAX over ranks of timings, SUM of counters, all-gather of the per-agent array (rank order =
    global agent id order).  Works on any backend (nccl on GPUs, gloo in the CPU tests)."""
    t = torch.tensor(timings, dtype=torch.float64, device=device)
    c = torch.tensor(counters, dtype=torch.float64, device=device)
    v = torch.as_tensor(per_agent, dtype=torch.float64).to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        out = [torch.empty_like(v) for _ in range(dist.get_world_size())]
        dist.all_gather(out, v)
        v = torch.cat(out)
    return t.tolist(), c.tolist(), v
