"""Dev helper (torchrun, N ranks): device->host bandwidth of a 35 MB pinned buffer per GPU,
all ranks copying at once, with and without NUMA binding; prints the topology on rank 0.
    python -m torch.distributed.run --nproc-per-node 8 scripts/d2h_topology.py"""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from ut_multirobot_sim_b200 import shard

rank = int(os.environ.get("RANK", "0")); lr = int(os.environ.get("LOCAL_RANK", "0"))
ws = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(lr)
if ws > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
if rank == 0:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
    print(subprocess.run(["bash", "-c", "lscpu | grep -i 'numa\\|socket\\|^CPU(s)'; nproc"], capture_output=True, text=True).stdout)
    sys.stdout.flush()
dev = torch.empty(35 << 20, dtype=torch.uint8, device="cuda")

def measure(tag):
    host = torch.empty(35 << 20, dtype=torch.uint8).pin_memory()
    host.zero_()
    for _ in range(3):
        host.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
    if ws > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        host.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = 50 * (35 << 20) / dt / 1e9
    t = torch.tensor([gbs], device="cuda", dtype=torch.float64)
    lst = [torch.zeros_like(t) for _ in range(ws)]
    if ws > 1:
        dist.all_gather(lst, t)
    else:
        lst = [t]
    if rank == 0:
        v = [float(x.item()) for x in lst]
        print("%s: per-GPU D2H GB/s %s  sum %.1f" % (tag, ["%.1f" % x for x in v], sum(v)))
        sys.stdout.flush()

measure("unbound, all ranks at once")
cpus = shard.bind_host_to_gpu(lr)
print("rank %d bound to %s" % (rank, None if cpus is None else "%d cpus %d..%d" % (len(cpus), cpus[0], cpus[-1])))
sys.stdout.flush()
measure("bound, all ranks at once")
if ws > 1:
    for r in range(ws):      # one rank at a time
        if r == rank:
            measure_alone = True
        dist.barrier()
    dist.destroy_process_group()
