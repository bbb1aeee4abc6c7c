"""Event-batch sharding of the sampling path over the GPUs of one node.

Showers are independent units (GroupNorm and attention are per-sample), so the batch is cut into contiguous
slices, one per rank, weights are replicated, there is no collective on the data path, and one all-gather
collects the showers at the end.  The per-step noise is keyed by the *global* shower index
(`sampler.shower_offset`), so the result does not depend on the number of ranks.
Works with any `torch.distributed` backend: `nccl` on the GPUs, `gloo` in the CPU tests.
"""
import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group; (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n_showers, rank, world_size):
    """Contiguous slice [lo, hi) of the event batch owned by `rank`; the first n % world ranks take one extra shower."""
    if n_showers < 0 or world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad shard request: n={n_showers} rank={rank} world={world_size}")
    base, extra = divmod(n_showers, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_showers(local, n_showers):
    """All-gather the ranks' shower slices (first dim) into the full batch, in global order, on every rank."""
    rank, ws = world()
    if ws == 1:
        return local
    sizes = [shard_range(n_showers, r, ws) for r in range(ws)]
    cap = max(hi - lo for lo, hi in sizes)
    width = 1
    for d in local.shape[1:]:
        width *= int(d)
    flat = local.reshape(local.shape[0], width)      # explicit width: an empty slice (n < world size) reshapes too
    if len({hi - lo for lo, hi in sizes}) == 1:
        out = torch.empty((ws * cap, flat.shape[1]), dtype=flat.dtype, device=flat.device)
        dist.all_gather_into_tensor(out, flat.contiguous())
    else:   # ragged split: pad every slice to the largest one, gather, drop the padding
        pad = torch.zeros((cap, flat.shape[1]), dtype=flat.dtype, device=flat.device)
        pad[: flat.shape[0]] = flat
        buf = torch.empty((ws * cap, flat.shape[1]), dtype=flat.dtype, device=flat.device)
        dist.all_gather_into_tensor(buf, pad)
        out = torch.cat([buf[r * cap: r * cap + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], dim=0)
    return out.reshape((n_showers,) + tuple(local.shape[1:]))


def sample_sharded(model, energy, layers, num_steps=400, sample_offset=0, start=None, seed=None):
    """`Diffusion.sample` for a batch that is split over the ranks: every rank passes the FULL conditioning
    (and optionally the full start noise), samples its own slice and receives the full set of showers.
    The start noise (when not passed in) and the per-step noise both come from the library's Philox stream keyed by
    GLOBAL shower index, so the showers do not depend on the number of ranks and no two ranks ever share a stream.
    `seed=None` draws the key from torch's global RNG: seed that identically on every rank (the per-step noise key is
    drawn from it as well).  Returns a device tensor (n_showers, *data_shape)."""
    rank, ws = world()
    n = energy.shape[0]
    lo, hi = shard_range(n, rank, ws)
    sampler = model.sampler_algorithm
    if start is None and seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())       # same draw on every (identically seeded) rank
    if hi == lo:
        # fewer showers than ranks: this rank owns none, but still consumes the RNG draw the others make and joins the gather
        sampler._seed()
        x = torch.empty([0] + list(model._data_shape), device=model.device, dtype=torch.float32)
        return gather_showers(x, n)
    prev = sampler.shower_offset
    sampler.shower_offset = lo
    try:
        e = energy[lo:hi]
        l = None if layers is None else layers[lo:hi]
        if start is None:
            start_l = model.philox_normal(hi - lo, seed, shower_offset=lo, slot=0x7fffffff)
        else:
            start_l = start[lo:hi]
        x, _, _ = sampler(model, start_l, e, l, num_steps, sample_offset, False)
    finally:
        sampler.shower_offset = prev
    return gather_showers(x, n)


# ----------------------------------------------------------------------------------------------------------------------
# Data-parallel training (BASELINE config 5): every rank runs the training step of
# train/train_diffusion.py:42-66 on its own batch, weights replicated; the only exchange is ONE all-reduce of the flat fp32
# gradient buffer (2.2 M elements, 8.85 MB for Dataset 2) between the backward pass and the optimizer step.  On NVSwitch
# that is tens of microseconds against a step of tens of milliseconds, so it is a single bucket enqueued behind the last
# backward kernel, not a pipeline of buckets.
# ----------------------------------------------------------------------------------------------------------------------
def allreduce_gradients(flat):
    """Sum the flat gradient buffer over the ranks in place; returns the divisor (world size) the optimizer step folds into
    its loss scale, so the average costs no extra pass over the buffer."""
    rank, ws = world()
    if ws > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    return float(ws)


def broadcast_parameters(module, src=0):
    """Rank `src`'s parameters and buffers to every rank (what DistributedDataParallel does at construction)."""
    rank, ws = world()
    if ws == 1:
        return
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.detach(), src)       # (detach shares the version counter: the engine sees the change)


def enable_data_parallel_training(model, src=0):
    """Make `model.training_loss` (and so `compute_loss(...).backward()` with any torch optimizer, or `fused_adam_step`) a
    data-parallel step: replicate rank `src`'s weights now, and all-reduce the flat gradient after every backward pass."""
    broadcast_parameters(model.model, src)
    model.grad_sync = allreduce_gradients
    return model


class DataParallelTrainer:
    """The reference's training loop body (train/train_diffusion.py:42-66: compute_loss -> backward -> optimizer.step) for one
    rank of a data-parallel job, all on the device: cd_train_step, one all-reduce of the flat gradient, fused Adam.

        trainer = DataParallelTrainer(model, lr=4e-4)
        for E, layers, data in loader:          # THIS rank's batch (`shard_range` cuts a global batch)
            loss = trainer.step(data, E, layers)
    """

    def __init__(self, model, lr=4e-4, betas=(0.9, 0.999), eps=1e-8, src=0):
        self.model, self.lr, self.betas, self.eps = enable_data_parallel_training(model, src), lr, betas, eps

    def step(self, data, E, layers=None, noise=None, time=None, rnd_normal=None, read_loss=True):
        """One optimizer step on this rank's batch.  Returns this rank's loss (a 0-d tensor), or None with read_loss=False
        (then nothing in the step synchronises with the host)."""
        m = self.model
        m._read_loss = read_loss
        try:
            loss = m.loss_function(m, data, E, noise=noise, time=time, layers=layers, rnd_normal=rnd_normal)
        finally:
            m._read_loss = True
        m.fused_adam_step(self.lr, self.betas, self.eps)
        return loss
