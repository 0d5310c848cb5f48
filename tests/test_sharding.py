"""N > 1 host logic on CPU: utterance sharding + host gather over a world_size-2 gloo group."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from open_universe_b200 import sharding


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 64, 4096):
        for w in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def test_length_buckets_are_exact_length():
    lengths = [100, 200, 100, 100, 200, 300]
    batches = sharding.length_buckets(lengths, range(6), 2)
    assert sorted(sum(batches, [])) == list(range(6))
    for b in batches:
        assert len({lengths[i] for i in b}) == 1 and len(b) <= 2


def _fake_enhance(x, gain=2.0):
    return x * gain + 1.0


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(0)
    clips = [torch.randn(50 + 10 * (i % 3), generator=g) for i in range(11)]
    out = sharding.enhance_sharded(_fake_enhance, clips, 4, rank, world, gain=3.0)
    if rank == 0:
        ok = all(torch.equal(o, c * 3.0 + 1.0) for o, c in zip(out, clips)) and len(out) == 11
        q.put(ok)
    else:
        assert out is None
    dist.destroy_process_group()


def test_enhance_sharded_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get() is True


def test_enhance_sharded_single_rank():
    clips = [torch.ones(5) * i for i in range(5)]
    out = sharding.enhance_sharded(_fake_enhance, clips, 2)
    assert all(torch.equal(o, c * 2.0 + 1.0) for o, c in zip(out, clips))
