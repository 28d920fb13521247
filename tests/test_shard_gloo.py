"""world_size-2 gloo test (CPU) of the multi-GPU host logic: contiguous stream shards, per-rank seeded
parameters/inputs, max-over-ranks timing, checksum gather.  The per-shard renders are done by the CPU
oracle here (test infrastructure); on the GPU box the same shard arithmetic feeds bench.py's ranks."""
import os
import socket

import numpy as np
import pytest

import za_oracle as zo
from za_gpu import shard
from za_gpu import workloads as wl


def test_shard_ranges_cover_and_balance():
    for n in (1, 7, 1024, 4096, 65536):
        for world in (1, 2, 4, 8):
            r = [shard.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:]))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_streams, nf, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a, b = shard.shard_range(n_streams, rank, world)
    p = wl.cascade8_params(b - a, first_stream=a)
    x = wl.noise_two_sines(b - a, nf, first_stream=a)
    y, _ = zo.PortPlugin("cascade8").render(p, x, threads=1)
    dist.barrier()
    t = shard.max_over_ranks(dist, 1.0 + rank)
    rows = shard.gather_to_rank0(dist, shard.stream_checksums(y), n_streams, world)
    if rank == 0:
        q.put((t, rows))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_matches_single_process():
    import torch.multiprocessing as mp
    n_streams, nf, world = 5, 4800, 2
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_streams, nf, q)) for r in range(world)]
    for p in procs:
        p.start()
    t, rows = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert t == 2.0                                              # max over ranks of (1.0, 2.0)
    y, _ = zo.PortPlugin("cascade8").render(wl.cascade8_params(n_streams), wl.noise_two_sines(n_streams, nf), threads=1)
    assert np.array_equal(rows, shard.stream_checksums(y))       # shards reproduce the single-process render bit for bit
