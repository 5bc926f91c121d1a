"""CPU, world_size 2 over gloo: the N > 1 host logic of bench.py -- frame
sharding by index with no data-path collective, the barrier, and the
max-over-ranks reduction of the timed region."""
import importlib.util
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sharding():
    spec = importlib.util.spec_from_file_location(
        "dc_sharding", os.path.join(ROOT, "depth-conversions_b200", "sharding.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = _sharding()
    F = 6
    mine = torch.tensor(list(sh.rank_frames(rank, world, F)), dtype=torch.int64)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)  # test-only: the product path never gathers
    dist.barrier()
    t = torch.tensor([1.5 if rank == 0 else 2.5], dtype=torch.float64)  # per-rank "device time"
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    blocks = sh.contiguous_blocks(0, 99, world)
    q.put((rank, [g.tolist() for g in gathered], float(t.item()), blocks))
    dist.destroy_process_group()


def test_two_ranks_shard_frames_disjointly():
    world, port = 2, 29533
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, gathered, tmax, blocks in results:
        frames = [f for g in gathered for f in g]
        assert sorted(frames) == list(range(12))          # every frame exactly once
        assert gathered[rank] == list(range(rank * 6, rank * 6 + 6))
        assert tmax == 2.5                                # max over ranks, not rank 0's clock
        assert blocks == [(0, 49), (50, 99)]


def test_contiguous_blocks_edge_cases():
    sh = _sharding()
    assert sh.contiguous_blocks(0, 9, 4) == [(0, 2), (3, 5), (6, 7), (8, 9)]
    # every part gets work: 9 frames on 8 GPUs use all 8
    assert sh.contiguous_blocks(0, 8, 8) == [(0, 1)] + [(i, i) for i in range(2, 9)]
    assert sh.contiguous_blocks(7, 3, 2) == [(3, 5), (6, 7)]      # reversed bounds
    assert sh.contiguous_blocks(5, 5, 8) == [(5, 5)]              # more GPUs than frames
    assert sh.contiguous_blocks(0, 99, 1) == [(0, 99)]
    for parts in range(1, 9):
        b = sh.contiguous_blocks(0, 9999, parts)
        assert b[0][0] == 0 and b[-1][1] == 9999
        assert all(b[i][1] + 1 == b[i + 1][0] for i in range(len(b) - 1))
        sizes = [l - f + 1 for f, l in b]
        assert len(b) == parts and max(sizes) - min(sizes) <= 1
