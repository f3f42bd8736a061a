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


def shard_worlds_weighted(total_worlds, rank, weights):
    """Contiguous block of rank `rank` when rank r should get a share proportional to weights[r]
    (e.g. its measured device->host link rate, for host-buffer stepping on a box whose GPUs do
    not all have the same path to host memory): returns (base, count).  Exact partition of
    [0, total_worlds): shares are floored and the remainder goes to the largest fractional
    parts (ties: lowest rank), so every rank computes the same table from the same weights.
    Worlds keep their global index (`world_index_base`), so the Philox streams -- every noise
    sample and pose -- are those of the unsharded job whatever the weights."""
    n = len(weights)
    if n <= 0 or not (0 <= rank < n) or total_worlds < 0:
        raise ValueError("bad shard request: total=%r rank=%r weights=%r" % (total_worlds, rank, weights))
    w = [float(x) if (x == x and x > 0.0) else 0.0 for x in weights]
    tot = sum(w)
    if tot <= 0.0:
        return shard_worlds(total_worlds, rank, n)
    exact = [total_worlds * x / tot for x in w]
    counts = [int(e) for e in exact]
    rest = total_worlds - sum(counts)
    order = sorted(range(n), key=lambda r: (-(exact[r] - counts[r]), r))
    for r in order[:rest]:
        counts[r] += 1
    return sum(counts[:rank]), counts[rank]


def exchange_rates(local_rate, group=None, device=None):
    """All-gather of one float per rank (a measured link rate) -> list indexed by rank.  Called once,
    before any timed region; the only other exchange of the design is reduce_stats."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [float(local_rate)]
    dev = device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu")
    t = torch.tensor([float(local_rate)], dtype=torch.float64, device=dev)
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size(group))]
    dist.all_gather(out, t, group=group)
    return [float(x.item()) for x in out]


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
