"""world_size-2 gloo run of the multi-GPU host logic: static instance
sharding, barrier, max-over-ranks timing reduction (bench.py's N>1 path)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from xrtailor_b200 import batch


def _worker(rank, world, port, n_inst, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = batch.shard(n_inst, world, rank)
    ids = torch.full((n_inst,), -1, dtype=torch.int64)
    ids[mine] = rank
    dist.all_reduce(ids, op=dist.ReduceOp.MAX)
    ms = torch.tensor([10.0 * (rank + 1)], dtype=torch.float64)  # rank 1 is the slow one
    dist.barrier()
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    frames = torch.tensor([len(mine) * 5], dtype=torch.int64)
    dist.all_reduce(frames, op=dist.ReduceOp.SUM)
    if rank == 0:
        out.put((ids.tolist(), float(ms.item()), int(frames.item())))
    dist.destroy_process_group()


def test_two_rank_sharding_and_timing():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    n_inst = 7
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_inst, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    owners, ms, frames = q.get()
    assert owners == [i % 2 for i in range(n_inst)]   # every instance owned exactly once, i mod G
    assert ms == 20.0                                   # max over ranks
    assert frames == n_inst * 5
    assert batch.frames_per_second(frames, [ms]) == n_inst * 5 / 0.02
