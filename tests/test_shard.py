"""Host-side multi-GPU logic on CPU: block partition + the statistics gather over
a world_size-2 gloo group (the only collective of the design)."""
import os
import socket

import numpy as np
import pytest

from ut_multirobot_sim_b200.shard import exchange_rates, reduce_stats, shard_worlds, shard_worlds_weighted


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


def test_weighted_shards_partition_exactly_and_follow_the_weights():
    # the 8-GPU box of this pool: 12.6 GB/s to GPUs 0-3, ~22 GB/s to GPUs 4-7 (profiles/r2_d2h_topology.txt)
    rates = [12.6, 12.6, 12.7, 12.6, 21.9, 22.8, 21.4, 22.1]
    for total in (0, 1, 7, 8191, 65536, 65537):
        blocks = [shard_worlds_weighted(total, r, rates) for r in range(8)]
        assert blocks[0][0] == 0
        for (b0, c0), (b1, _) in zip(blocks, blocks[1:]):
            assert b0 + c0 == b1
        assert blocks[-1][0] + blocks[-1][1] == total
        for (_, c), w in zip(blocks, rates):
            assert abs(c - total * w / sum(rates)) < 1.0 + 1e-9
    counts = [shard_worlds_weighted(65536, r, rates)[1] for r in range(8)]
    # per-rank copy time is equalised: worlds / rate within one world of each other
    t = [c / w for c, w in zip(counts, rates)]
    assert max(t) - min(t) <= 1.0 / min(rates) + 1e-9
    # equal weights == the uniform partition; useless weights fall back to it
    for r in range(4):
        assert shard_worlds_weighted(1001, r, [3.0] * 4) == shard_worlds(1001, r, 4)
        assert shard_worlds_weighted(1001, r, [0.0, float("nan"), -1.0, 0.0]) == shard_worlds(1001, r, 4)
    assert shard_worlds_weighted(10, 1, [1.0, 0.0, 1.0]) == (5, 0)      # a dead link gets nothing
    with pytest.raises(ValueError):
        shard_worlds_weighted(10, 3, [1.0, 1.0])
    assert exchange_rates(7.5) == [7.5]                                  # no process group: identity


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
    # link rates are exchanged once; every rank derives the same weighted table from them
    rates = exchange_rates([10.0, 30.0][rank])
    wb, wc = shard_worlds_weighted(total, rank, rates)
    ok_w = rates == [10.0, 30.0] and (wb, wc) == ((0, 250) if rank == 0 else (250, 751))
    mine_w = ranges[wb:wb + wc]
    gw = reduce_stats({"hits": float(len(mine_w)), "min": float(mine_w.min()), "max": float(mine_w.max()),
                       "sum": float(mine_w.sum())})
    ok_w = ok_w and gw["hits"] == total and abs(gw["sum"] - ranges.sum()) < 1e-6
    g = reduce_stats(local)
    ok = (g["hits"] == total and abs(g["sum"] - ranges.sum()) < 1e-6 and g["min"] == ranges.min()
          and g["max"] == ranges.max())
    # a rank with no hits must not poison the minimum
    g2 = reduce_stats({"hits": 0.0, "min": 0.0, "max": 0.0, "sum": 0.0} if rank == 0 else local)
    ok = ok and g2["min"] == float(mine.min()) if rank == 1 else ok
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("1" if (ok and ok_w) else "0")


def test_stats_gather_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        assert open(tmp_path / ("ok%d" % r)).read() == "1"


def test_reduce_stats_without_process_group_is_identity():
    s = {"hits": 3.0, "min": 1.0, "max": 2.0, "sum": 4.5}
    assert reduce_stats(s) == s
