"""world_size-2 gloo tests of the event-batch sharding: the gathered result equals the single-process one."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from calodiffusion_b200 import distributed as D


def test_shard_range_partitions_the_batch():
    for n in (0, 1, 7, 8, 1000, 1001):
        for ws in (1, 2, 3, 8):
            spans = [D.shard_range(n, r, ws) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(ws - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        D.shard_range(8, 2, 2)


class _FakeSampler:
    """Stands in for a sampler: the 'shower' of global index i is a deterministic function of i, its conditioning and start."""
    shower_offset = 0

    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        idx = torch.arange(start.shape[0], dtype=torch.float32) + self.shower_offset
        x = start * 0.5 + energy.view(-1, 1, 1) + idx.view(-1, 1, 1) * 10.0 + (0 if layers is None else layers.sum(1).view(-1, 1, 1))
        return x, [], []


class _FakeModel:
    _data_shape = [3, 4]

    def __init__(self):
        self.sampler_algorithm = _FakeSampler()

    def noise_generation(self, shape):
        raise AssertionError("the test passes the start noise explicitly")


def _expected(E, L, start):
    m = _FakeModel()
    x, _, _ = m.sampler_algorithm(m, start, E, L, 4, 0, False)
    return x


def _worker(rank, ws, port, n, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        g = torch.Generator().manual_seed(5)
        E, L, start = torch.rand(n, 1, generator=g), torch.randn(n, 6, generator=g), torch.randn(n, 3, 4, generator=g)
        m = _FakeModel()
        out = D.sample_sharded(m, E, L, num_steps=4, start=start)
        assert m.sampler_algorithm.shower_offset == 0          # restored
        ok = torch.equal(out, _expected(E, L, start))
        q.put((rank, bool(ok), tuple(out.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [8, 7])   # even and ragged split
def test_sample_sharded_world_size_2_gloo(n):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), res
    assert all(r[2] == (n, 3, 4) for r in res)
