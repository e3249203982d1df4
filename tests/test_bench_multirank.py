"""The N>1 path of bench.py on CPU: two gloo ranks, max-over-ranks timing and whole-job aggregation
(replicas only — there is no data-path collective to test)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import bench


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # rank r pretends to have needed (100 + 50 r) ms of decode time and so on
    vals = bench.max_over_ranks(dist, [100.0 + 50.0 * rank, 1.0 + rank, 17.0 - rank, 0.01 * (rank + 1)], "cpu")
    if rank == 0:
        out.put(vals)
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_max_and_aggregate():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    vals = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert vals == [150.0, 2.0, 17.0, 0.02]
    # 2 ranks x 3 requests x 127 decoded tokens in 150 ms
    assert abs(bench.aggregate_tok_per_s(2, 3, 128, vals[0]) - 2 * 3 * 127 / 0.150) < 1e-6
    assert bench.max_over_ranks(None, [1.0, 2.0], "cpu") == [1.0, 2.0]


def test_synthetic_prompt_is_deterministic_and_in_range():
    a, b = bench.synthetic_prompt(128256, 512), bench.synthetic_prompt(128256, 512)
    assert (a == b).all() and a.min() >= 0 and a.max() < 128000 and a.size == 512
