"""Multi-GPU sharding of batched worlds.

Worlds are independent (nothing is shared between them; the static map is
read-only and replicated per GPU), so a job of W worlds is cut into contiguous
blocks, one per rank, and every rank steps its block with NO collective on the
step path.  The only exchange is the final statistics gather (K3's four
numbers per rank), done with torch.distributed (NCCL on GPUs, gloo in the CPU
tests).  `world_index_base` keeps the Philox counters global, so a sharded run
produces exactly the ranges of the unsharded one.
"""


def shard_worlds(total_worlds, rank, world_size):
    """Contiguous block of rank `rank`: returns (base, count).  The remainder is
    spread over the first ranks, so counts differ by at most one."""
    if world_size <= 0 or not (0 <= rank < world_size) or total_worlds < 0:
        raise ValueError("bad shard request: total=%r rank=%r world_size=%r" % (total_worlds, rank, world_size))
    q, r = divmod(total_worlds, world_size)
    count = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, count


def reduce_stats(stats, group=None, device=None):
    """All-reduce of World.stats() dicts {hits, min, max, sum} over the ranks."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(stats)
    dev = device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu")
    s = torch.tensor([stats["hits"], stats["sum"]], dtype=torch.float64, device=dev)
    # ranks without a single hit report min 0: keep them out of the global minimum
    mn = torch.tensor([stats["min"] if stats["hits"] > 0 else float("inf")], dtype=torch.float64, device=dev)
    mx = torch.tensor([stats["max"]], dtype=torch.float64, device=dev)
    dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    hits = float(s[0].item())
    return {"hits": hits, "sum": float(s[1].item()),
            "min": float(mn.item()) if hits > 0 else 0.0, "max": float(mx.item())}


def bind_host_to_gpu(device_index):
    """Pins the calling process to the CPU cores next to GPU `device_index` (NVML's ideal
    affinity), so that the pinned host buffers a rank allocates afterwards -- first touch --
    live on the GPU's own NUMA node and its device->host copies do not cross the socket
    link.  Matters for the host-buffer step at 8 GPUs (35 MB of ranges per GPU per step).
    Returns the CPU list, or None when NVML / the affinity call is unavailable (no error:
    the binding is an optimisation)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        # honour CUDA_VISIBLE_DEVICES remapping
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = device_index
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if device_index < len(ids) and ids[device_index].isdigit():
                idx = int(ids[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [64 * i + b for i, word in enumerate(mask) for b in range(64) if (word >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None
