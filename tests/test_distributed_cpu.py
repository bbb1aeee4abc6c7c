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

    def _seed(self):
        return int(torch.randint(0, 2 ** 62, (1,)).item())

    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        idx = torch.arange(start.shape[0], dtype=torch.float32) + self.shower_offset
        x = start * 0.5 + energy.view(-1, 1, 1) + idx.view(-1, 1, 1) * 10.0 + (0 if layers is None else layers.sum(1).view(-1, 1, 1))
        return x, [], []


class _FakeModel:
    _data_shape = [3, 4]
    device = torch.device("cpu")

    def __init__(self):
        self.sampler_algorithm = _FakeSampler()

    def philox_normal(self, n_showers, seed, shower_offset=0, slot=0):
        """Stand-in for the library's Philox stream: row i is a function of (seed, global shower index) only."""
        rows = [torch.randn(3, 4, generator=torch.Generator().manual_seed((int(seed) + 7919 * (shower_offset + i)) % (2 ** 31)))
                for i in range(int(n_showers))]
        return torch.stack(rows) if rows else torch.empty(0, 3, 4)

    def noise_generation(self, shape):
        raise AssertionError("the test passes the start noise explicitly")


def _expected(E, L, start):
    m = _FakeModel()
    x, _, _ = m.sampler_algorithm(m, start, E, L, 4, 0, False)
    return x


def _worker(rank, ws, port, n, q, own_start=True):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        g = torch.Generator().manual_seed(5)
        E, L, start = torch.rand(n, 1, generator=g), torch.randn(n, 6, generator=g), torch.randn(n, 3, 4, generator=g)
        m = _FakeModel()
        if own_start:
            out = D.sample_sharded(m, E, L, num_steps=4, start=start)
        else:
            # start=None, seed=None: the key is drawn from the identically seeded torch RNG and the start noise is keyed by
            # global shower index - the gathered result must equal the single-process one (no duplicated streams)
            torch.manual_seed(123)
            out = D.sample_sharded(m, E, L, num_steps=4)
            torch.manual_seed(123)
            start = m.philox_normal(n, int(torch.randint(0, 2 ** 62, (1,)).item()))
            rows = out - _expected(E, L, torch.zeros_like(start))     # = 0.5 * start row by row
            assert n < 2 or not torch.equal(rows[0], rows[n - 1])     # distinct ranks do not reuse a stream
        assert m.sampler_algorithm.shower_offset == 0          # restored
        ok = torch.allclose(out, _expected(E, L, start), atol=0, rtol=0) if own_start else torch.allclose(out, _expected(E, L, start))
        q.put((rank, bool(ok), tuple(out.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,own_start", [(8, True), (7, True), (7, False), (1, False)])   # even / ragged / fewer showers than ranks
def test_sample_sharded_world_size_2_gloo(n, own_start):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q, own_start)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), res
    assert all(r[2] == (n, 3, 4) for r in res)


# ---- data-parallel training (row E2): one all-reduce of the flat gradient between backward and the optimizer step ----------
class _FakeTrainModel:
    """Stands in for CaloDiffusion's training interface with a linear least-squares model in torch: the same hooks
    (`grad_sync`, flat gradient buffer times a loss scale, fused Adam dividing by loss_scale * grad_div)."""

    def __init__(self, seed):
        self.model = torch.nn.Linear(5, 1, bias=True)
        with torch.no_grad():
            for p in self.model.parameters():
                p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(seed)))
        self.grad_sync = None
        self.G = torch.zeros(6)
        self.m, self.v, self.t = torch.zeros(6), torch.zeros(6), 0
        self.loss_scale, self.grad_div = 64.0, 1.0

    def loss_function(self, model, data, E, noise=None, time=None, layers=None, rnd_normal=None):
        w, b = self.model.weight.detach().reshape(-1), self.model.bias.detach()
        r = E @ w + b - data
        g = torch.cat([(2 * r.view(-1, 1) * E).mean(0), (2 * r).mean().view(1)])
        self.G = g * self.loss_scale
        self.grad_div = float(self.grad_sync(self.G)) if self.grad_sync is not None else 1.0
        return (r ** 2).mean() if getattr(self, "_read_loss", True) else None

    def fused_adam_step(self, lr, betas, eps):
        g = self.G / (self.loss_scale * self.grad_div)
        self.t += 1
        self.m = betas[0] * self.m + (1 - betas[0]) * g
        self.v = betas[1] * self.v + (1 - betas[1]) * g * g
        upd = lr * (self.m / (1 - betas[0] ** self.t)) / ((self.v / (1 - betas[1] ** self.t)).sqrt() + eps)
        with torch.no_grad():
            self.model.weight -= upd[:5].view(1, 5)
            self.model.bias -= upd[5:]


def _train_data(n):
    g = torch.Generator().manual_seed(3)
    return torch.randn(n, 5, generator=g), torch.randn(n, generator=g)


def _train_worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        m = _FakeTrainModel(seed=100 + rank)                 # ranks start from DIFFERENT weights: the broadcast must fix that
        tr = D.DataParallelTrainer(m, lr=1e-2)
        X, y = _train_data(16)
        lo, hi = D.shard_range(16, rank, ws)
        losses = [float(tr.step(y[lo:hi], X[lo:hi])) for _ in range(3)]
        assert tr.step(y[lo:hi], X[lo:hi], read_loss=False) is None
        q.put((rank, m.model.weight.detach().clone(), m.model.bias.detach().clone(), losses, m.grad_div))
    finally:
        dist.destroy_process_group()


def test_data_parallel_training_world_size_2_gloo():
    """Two ranks on half a batch each end up with identical weights, equal to one process stepping on the full batch (equal
    shard sizes: the mean of the shard gradients is the full-batch gradient), starting from rank 0's weights."""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_train_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert torch.equal(res[0][1], res[1][1]) and torch.equal(res[0][2], res[1][2])
    assert res[0][4] == 2.0 and res[1][4] == 2.0
    single = _FakeTrainModel(seed=100)
    tr = D.DataParallelTrainer(single, lr=1e-2)              # no process group: world size 1, no collective
    X, y = _train_data(16)
    for _ in range(4):
        tr.step(y, X)
    assert single.grad_div == 1.0
    assert torch.allclose(single.model.weight, res[0][1], atol=1e-6), (single.model.weight, res[0][1])
    assert torch.allclose(single.model.bias, res[0][2], atol=1e-6)
