"""Multi-GPU host logic on the CPU: work items are dealt to ranks without
any data-path collective; two gloo processes each code their share (with the oracle -- no GPU
here) and the job's per-item table and step time come together over torch.distributed."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from flashgmm_b200 import shard, synth


def test_tile_grid_of_a_4k_image():
    tiles = shard.tile_image(2160, 3840)
    assert len(tiles) == 25                                   # 5 x 5, bottom padded to 2560 (BASELINE.json configs[4])
    assert tiles[0] == shard.Tile(0, 512, 0, 768, 0, 0)
    assert tiles[-1] == shard.Tile(2048, 2560, 3072, 3840, 400, 0)
    assert sum(1 for t in tiles if t.pad_h) == 5 and not any(t.pad_w for t in tiles)
    assert shard.tile_image(512, 768) == [shard.Tile(0, 512, 0, 768, 0, 0)]
    assert len(shard.tile_image(513, 769)) == 4
    with pytest.raises(ValueError):
        shard.tile_image(0, 10)


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_assignment_covers_every_item_once_and_is_balanced(world):
    costs = [1.0] * 24 + [1.0] * 25                            # 24 images + 25 tiles: 49 equal items
    table = shard.assign(costs, world)
    flat = sorted(i for lst in table for i in lst)
    assert flat == list(range(49))
    sizes = [len(lst) for lst in table]
    assert max(sizes) - min(sizes) <= 1
    assert table == shard.assign(costs, world)                 # deterministic: no communication needed
    for r in range(world):
        assert shard.my_items(costs, world, r) == table[r]


def test_assignment_by_cost_puts_the_long_items_apart():
    costs = [10, 1, 1, 1, 10, 1, 1, 1, 1, 1]
    a, b = shard.assign(costs, 2)
    assert (0 in a) != (4 in a)                                 # the two expensive items on different ranks
    assert abs(sum(costs[i] for i in a) - sum(costs[i] for i in b)) <= 1
    with pytest.raises(ValueError):
        shard.my_items(costs, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _item(i):
    return synth.gmm_numpy(400 + 37 * i, seed=5000 + i)


def _worker(rank, world, port, n_items, out_dir):
    from oracle import port as oracle
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        costs = [400 + 37 * i for i in range(n_items)]
        mine = shard.my_items(costs, world, rank)
        coded = []
        for i in mine:
            sym, sg, mu, w = _item(i)
            b = oracle.gmm_encode(sym, sg, mu, w, 0)
            assert np.array_equal(oracle.gmm_decode(b, sg, mu, w, synth.search_radius(sym), 0), sym)
            coded.append((i, len(b)))
        table = shard.gather_item_sizes(n_items, coded)
        step = shard.max_over_ranks(0.25 + rank)                # rank 1 is the slow one
        torch.save({"mine": mine, "table": table, "step": step}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_two_ranks_code_their_shares_and_agree_on_the_job(tmp_path):
    from oracle import port as oracle
    world, n_items = 2, 7
    mp.spawn(_worker, args=(world, _free_port(), n_items, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    assert sorted(res[0]["mine"] + res[1]["mine"]) == list(range(n_items))
    assert not set(res[0]["mine"]) & set(res[1]["mine"])
    want = [len(oracle.gmm_encode(*_item(i), 0)) for i in range(n_items)]     # sharding changes no byte
    assert res[0]["table"] == want and res[1]["table"] == want
    assert res[0]["step"] == res[1]["step"] == 1.25              # max over ranks


def test_single_process_needs_no_process_group():
    assert shard.max_over_ranks(0.5) == 0.5
    assert shard.gather_item_sizes(3, [(0, 10), (2, 30)]) == [10, -1, 30]


def test_sharded49_job_is_dealt_evenly():
    """BASELINE.json configs[4] as bench.py --workload sharded49 builds it: 24 images + a 3840x2160 image as 5 x 5 tiles of
    768x512 = 49 equal-cost items; every rank count of one box gets them all exactly once, loads differ by at most one."""
    tiles = shard.tile_image(2160, 3840)
    assert len(tiles) == 25 and tiles[-1].pad_h == 2560 - 2160 and all(t.pad_w == 0 for t in tiles)
    costs = [1.0] * (24 + len(tiles))
    for world in (1, 2, 4, 8):
        parts = shard.assign(costs, world)
        assert sorted(i for p in parts for i in p) == list(range(49))
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1
        assert [shard.my_items(costs, world, r) for r in range(world)] == parts
