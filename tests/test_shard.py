"""Host-side multi-GPU logic on CPU: block partition + the statistics gather over
a world_size-2 gloo group (the only collective of the design)."""
import os
import socket

import numpy as np
import pytest

from ut_multirobot_sim_b200.shard import reduce_stats, shard_worlds


def test_shard_worlds_partitions_exactly():
    for total in (0, 1, 7, 4096, 65536, 65537):
        for n in (1, 2, 3, 4, 8):
            blocks = [shard_worlds(total, r, n) for r in range(n)]
            assert blocks[0][0] == 0
            for (b0, c0), (b1, _) in zip(blocks, blocks[1:]):
                assert b0 + c0 == b1
            assert blocks[-1][0] + blocks[-1][1] == total
            counts = [c for _, c in blocks]
            assert max(counts) - min(counts) <= 1
    assert shard_worlds(65536, 3, 8) == (3 * 8192, 8192)
    with pytest.raises(ValueError):
        shard_worlds(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    total = 1001
    base, count = shard_worlds(total, rank, world_size)
    # stand-in for the per-rank K3 result: "ranges" of this rank's worlds
    rng = np.random.default_rng(5)
    ranges = rng.uniform(0.5, 30.0, total)
    mine = ranges[base:base + count]
    local = {"hits": float(len(mine)), "min": float(mine.min()), "max": float(mine.max()),
             "sum": float(mine.sum())}
    if rank == 1:
        pass
    g = reduce_stats(local)
    ok = (g["hits"] == total and abs(g["sum"] - ranges.sum()) < 1e-6 and g["min"] == ranges.min()
          and g["max"] == ranges.max())
    # a rank with no hits must not poison the minimum
    g2 = reduce_stats({"hits": 0.0, "min": 0.0, "max": 0.0, "sum": 0.0} if rank == 0 else local)
    ok = ok and g2["min"] == float(mine.min()) if rank == 1 else ok
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("1" if ok else "0")


def test_stats_gather_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        assert open(tmp_path / ("ok%d" % r)).read() == "1"


def test_reduce_stats_without_process_group_is_identity():
    s = {"hits": 3.0, "min": 1.0, "max": 2.0, "sum": 4.5}
    assert reduce_stats(s) == s
