"""Multi-GPU host logic on CPU: world_size-2 gloo process group.  The per-patch map is a
stand-in (a deterministic torch function of each patch on its own) -- what is under test is the partition, the padded
all_gather and the re-assembly order, which are backend independent."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from puflow_b200.sharding import block_range, patch_sharded, shard_files


def test_block_range_partitions_exactly():
    for n in (0, 1, 7, 32, 78, 4096):
        for world in (1, 2, 3, 4, 8):
            blocks = [block_range(n, r, world) for r in range(world)]
            covered = [i for lo, hi in blocks for i in range(lo, hi)]
            assert covered == list(range(n)), (n, world, blocks)
            assert max(hi - lo for lo, hi in blocks) == -(-n // world)
    assert [block_range(78, r, 8) for r in (0, 6, 7)] == [(0, 10), (60, 70), (70, 78)]
    with pytest.raises(ValueError):
        block_range(10, 2, 2)


def test_shard_files_round_robin():
    paths = [f"d/{c}.xyz" for c in "hgfedcba"]
    for world in (2, 3, 8, 16):
        got = [shard_files(paths, r, world) for r in range(world)]
        assert sorted(p for g in got for p in g) == sorted(paths)
        assert all(len(g) <= -(-len(paths) // world) for g in got)
    assert shard_files(paths, 0, 1) == paths      # single process: the reference's own order (upsample.py:38)


def _fake_sample(p: torch.Tensor) -> torch.Tensor:
    """[b, N, 3] -> [b, 4N, 3]: depends on each patch alone (like PointInterpFlow.sample in eval mode)."""
    m = p.mean(1, keepdim=True)
    return torch.cat([p * (r + 1) + m * r for r in range(4)], 1)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, Bp, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(11)
        patches = torch.randn(Bp, 16, 3, generator=g)            # identical on every rank
        calls = []

        def fn(p):
            calls.append(p.shape[0])
            return _fake_sample(p)

        got = patch_sharded(fn, patches)
        ok = torch.equal(got, _fake_sample(patches))
        lo, hi = block_range(Bp, rank, world)
        ok = ok and calls == [max(hi - lo, 1)]
        # max-over-ranks timing reduction as bench.py does it
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = ok and t.item() == world
        torch.save({"ok": bool(ok), "calls": calls}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("Bp", [78, 1, 7])
def test_patch_sharded_gloo_world2(tmp_path, Bp):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, Bp, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(tmp_path / f"r{r}.pt")
        assert res["ok"], (r, res)


def test_patch_sharded_without_group_is_identity_wrapper():
    p = torch.randn(5, 8, 3)
    assert torch.equal(patch_sharded(_fake_sample, p), _fake_sample(p))
