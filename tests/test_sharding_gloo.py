"""World-size-2 gloo test of the multi-rank host logic (runs on CPU): shard ranges partition the
global env ids, the statistics all-reduce sums over ranks, timing is the max over ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sparc_b200.sharding import all_reduce_stats, max_over_ranks, shard_range

    first, n_local = shard_range(total, world, rank)
    ids = torch.arange(first, first + n_local, dtype=torch.float64)
    # a fake per-rank statistics vector whose global sum is known in closed form
    stats = torch.tensor([ids.sum().item(), float(n_local), 1.0, 0, 0, n_local * 1000.0, 0, 0], dtype=torch.float64)
    all_reduce_stats(stats)
    t = max_over_ranks(0.5 + rank)
    torch.save({"first": first, "n": n_local, "stats": stats, "t": t}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [4096, 4097, 7])
def test_two_rank_sharding_and_stats_allreduce(tmp_path, total):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    assert res[0]["first"] == 0 and res[1]["first"] == res[0]["n"]
    assert res[0]["n"] + res[1]["n"] == total and abs(res[0]["n"] - res[1]["n"]) <= 1
    for r in res:
        assert r["stats"][0].item() == total * (total - 1) / 2      # sum of all global env ids
        assert r["stats"][1].item() == total and r["stats"][2].item() == world
        assert r["stats"][5].item() == total * 1000.0
        assert r["t"] == 1.5                                         # max over ranks


def test_shard_range_covers_everything():
    from sparc_b200.sharding import shard_range

    for total in (1, 8, 65536, 65537):
        for world in (1, 2, 4, 8):
            spans = [shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0
            for (a, n), (b, _) in zip(spans, spans[1:]):
                assert a + n == b
            assert spans[-1][0] + spans[-1][1] == total
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)
