"""north_star: "NCCL used only to gather outputs" -- sharding.gather_variable (all_gather of lengths + ncclGather of the
padded waveforms to rank 0) on the NCCL backend with device tensors, world size 2.  Needs two GPUs (skipped otherwise; the
host logic is covered on gloo by tests/test_sharding_gloo.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from moss_tts_b200.sharding import gather_variable, shard_range

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    n = 7
    b, e = shard_range(n, world, rank)
    rows = [torch.full((1000 + 37 * i,), float(i), device=dev) for i in range(b, e)]      # ragged "waveforms" on the device
    out = gather_variable(rows, device=dev)
    if rank == 0:
        ok = len(out) == n and all(o.is_cuda and o.numel() == 1000 + 37 * i and bool((o == i).all()) for i, o in enumerate(out))
        q.put(ok)
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_gather_outputs_world2_nccl():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
