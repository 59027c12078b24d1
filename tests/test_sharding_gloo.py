"""N>1 host logic on CPU: world_size-2 gloo — shard ranges cover the batch, outputs gather in order."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from moss_tts_b200.sharding import shard_range, gather_variable


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 129):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and 0 <= (b - a) - (d - c) <= 1


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 7
    b, e = shard_range(n, world, rank)
    # each request i yields a "waveform" of length 10+i filled with i
    rows = [torch.full((10 + i,), float(i)) for i in range(b, e)]
    out = gather_variable(rows)
    if rank == 0:
        ok = len(out) == n and all(o.numel() == 10 + i and bool((o == i).all()) for i, o in enumerate(out))
        q.put(ok)
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


def test_gather_outputs_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
