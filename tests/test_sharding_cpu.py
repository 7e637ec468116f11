"""CPU, world_size 2, gloo: the pair-sharding host logic.  The per-rank
'forward' is a stand-in that tags each pair, so the test checks partitioning + final gather
only -- it never routes the product path through a CPU implementation."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import pkg


def _worker(rank, world, port, total, n, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S = pkg("sharding")
    data = {"keypoints0": torch.arange(total, dtype=torch.float32)[:, None, None].expand(total, n, 2).contiguous(),
            "view0": {"image_size": torch.ones(total, 2)}, "scalar": 3}
    mine = S.shard_batch(data, world, rank)
    lo, hi = S.shard_range(total, world, rank)
    assert mine["keypoints0"].shape[0] == hi - lo and mine["view0"]["image_size"].shape[0] == hi - lo
    assert mine["scalar"] == 3
    ids = mine["keypoints0"][:, :, 0]  # pair id broadcast over keypoints
    pred = {"matches0": ids.long(), "matches1": (ids * 2).long(), "matching_scores0": ids / 10, "matching_scores1": ids / 20}
    full = S.gather_results(pred, total)
    want = torch.arange(total, dtype=torch.float32)[:, None].expand(total, n)
    ok = (torch.equal(full["matches0"], want.long()) and torch.equal(full["matches1"], (want * 2).long())
          and torch.allclose(full["matching_scores0"], want / 10) and full["matches0"].shape == (total, n))
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_shard_and_gather_world2():
    for total in (8, 5):  # even split, ragged tail
        ret = mp.Manager().dict()
        port = 29500 + os.getpid() % 2000 + total
        mp.spawn(_worker, args=(2, port, total, 6, ret), nprocs=2, join=True)
        assert ret[0] and ret[1], total


def test_shard_range_covers_everything():
    S = pkg("sharding")
    for total in (1, 7, 32, 1000):
        for world in (1, 2, 4, 8):
            spans = [S.shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
