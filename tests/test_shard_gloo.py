"""CPU, world_size 2 over gloo: the host-side sharding logic of the multi-GPU path (shard_range,
ragged all-gather, compute_sharded with a stand-in matcher -- the CUDA matcher needs a GPU)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stereovision_b200.shard import compute_sharded, gather_outputs, shard_range, shard_sizes


def test_shard_range_partitions():
    for n in (0, 1, 5, 8, 31, 512):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sz = shard_sizes(n, world)
            assert sum(sz) == n and max(sz) - min(sz) <= 1


class _FakeMatcher:
    """Stands in for StereoSGBM on CPU: a deterministic per-pair function of the inputs."""
    def compute(self, l, r):
        return (l.to(torch.int16) * 3 - r.to(torch.int16)).contiguous()


def _worker(rank, world, port, B, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(5)
        L = torch.randint(0, 256, (B, 6, 9), dtype=torch.uint8, generator=g)
        R = torch.randint(0, 256, (B, 6, 9), dtype=torch.uint8, generator=g)
        a, b = shard_range(B)
        local = compute_sharded(_FakeMatcher(), L, R, gather=False)
        assert local.shape[0] == b - a
        full = compute_sharded(_FakeMatcher(), L, R, gather=True)
        ref = _FakeMatcher().compute(L, R)
        ok = bool(torch.equal(full, ref)) and bool(torch.equal(gather_outputs(local, B), ref))
        q.put((rank, ok, a, b))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(B):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=60) for _ in ps]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    return sorted(res)


def test_two_ranks_even_and_ragged():
    for B in (8, 5, 1):
        res = _run(B)
        assert all(ok for _, ok, _, _ in res)
        assert res[0][2] == 0 and res[0][3] == res[1][2] and res[1][3] == B
