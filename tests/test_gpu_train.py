"""Training step on the device (BASELINE config 5) against the CPU oracle's autograd (itself pinned to the live
reference's compute_loss().backward() by tests/test_train_oracle.py): the loss, the gradient of every parameter tensor, one Adam
update, and the drop-in form `loss.backward(); torch.optim.Adam.step()`.
Gate: relative L2 per parameter tensor (north_star's 16-bit bound 1e-2; activations and activation gradients are fp16)."""
import os

import numpy as np
import pytest
import torch

from conftest import load_case, rel_l2
from test_train_oracle import train_inputs

pytestmark = pytest.mark.gpu


def _model(case, max_batch=2):
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case(case)
    cfg = dict(c["cfg"], SAMPLER="DDPM")
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=max_batch)
    m.load_state_dict(dict(c["sd"]))
    return m, c


def _oracle(c, E, layers, data, noise, rnd, time):
    from oracle.train_oracle import loss_and_grads, sigma_for
    torch.set_num_threads(os.cpu_count() or 1)
    return loss_and_grads(c["cfg"], c["sd"], data, E, layers, noise, sigma_for(c["cfg"], rnd, time))


@pytest.mark.parametrize("case", ["ds2", "ds3", "hgcal"])
def test_gradients_match_oracle(case):
    m, c = _model(case)
    E, layers, data, noise, rnd, time = train_inputs(c["shape"], 2, c["meta"]["cond_e"], c["meta"]["cond_layers"])
    loss_ref, grads_ref = _oracle(c, E, layers, data, noise, rnd, time)
    Lc = None if layers is None else layers.cuda()
    loss = m.loss_function(m, data.cuda(), E.cuda(), noise=noise.cuda(), time=time.cuda(), layers=Lc, rnd_normal=rnd.cuda())
    loss.backward()
    e_loss = abs(float(loss) - loss_ref) / abs(loss_ref)
    errs = {}
    for name, p in m.model.named_parameters():
        g = grads_ref["model." + name].numpy()
        assert p.grad is not None, name
        errs[name] = rel_l2(p.grad.cpu().numpy(), g)
    worst = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    total = rel_l2(np.concatenate([p.grad.cpu().numpy().ravel() for _, p in m.model.named_parameters()]),
                   np.concatenate([grads_ref["model." + n].numpy().ravel() for n, _ in m.model.named_parameters()]))
    print(f"{case}: loss rel.err {e_loss:.2e}; all-parameter gradient rel-L2 {total:.2e}; worst tensors: " +
          ", ".join(f"{k} {v:.2e}" for k, v in worst))
    # measured on B200: ds2 total 2.5e-3, worst tensor 6.1e-3; ds3 total 3.4e-4, worst 2.2e-3; hgcal shape (96-channel level,
    # PHI_INPUT, 32 conditioning inputs) total 4.7e-4, worst 3.4e-3; loss 5e-5 (hgcal 1.3e-4)
    assert e_loss < 5e-4
    assert total < {"ds2": 6e-3, "ds3": 1.5e-3, "hgcal": 1.5e-3}[case]      # measured 2.6e-3 / 3.4e-4 / 4.7e-4
    # worst single tensor over ten runs of this round: 6.1e-3 ... 7.7e-3 (ds2, a 32-element bias deep in the network; the
    # accumulation order of the atomics varies run to run), 2.4e-3 (ds3): gate at 2x the largest measured
    assert max(errs.values()) < 1.5e-2, worst


def test_fused_adam_matches_torch_adam():
    """One optimizer step: cd_train_adam on the device-resident master parameters vs torch.optim.Adam on the oracle's gradients."""
    m, c = _model("ds2")
    E, layers, data, noise, rnd, time = train_inputs(c["shape"], 2, 1, 46)
    _, grads_ref = _oracle(c, E, layers, data, noise, rnd, time)
    ref = {k: v.clone().requires_grad_(True) for k, v in c["sd"].items()}
    opt = torch.optim.Adam(list(ref.values()), lr=4e-4)
    for k, v in ref.items():
        v.grad = grads_ref[k]
    opt.step()
    m.compute_loss(data.cuda(), E.cuda(), noise.cuda(), layers.cuda(), rnd_normal=rnd.cuda())
    m.fused_adam_step(lr=4e-4)
    sd = m.state_dict()                      # pulls the updated master copy back into the torch parameters
    # Adam's first step is lr * g / (|g| + eps): where |g| >> eps it is lr*sign(g) and any gradient with the right sign matches,
    # where |g| ~ eps (1e-8) it amplifies the gradient's relative error.  Compare the update over all parameters at once, and
    # element-wise on the entries whose reference gradient is well above eps.
    steps, steps_ref, big = [], [], []
    for k, v in ref.items():
        steps_ref.append((v.detach() - c["sd"][k]).numpy().ravel())
        steps.append((sd[k].cpu() - c["sd"][k]).numpy().ravel())
        big.append(np.abs(grads_ref[k].numpy().ravel()) > 1e-6)
    steps, steps_ref, big = np.concatenate(steps), np.concatenate(steps_ref), np.concatenate(big)
    e_all, e_big = rel_l2(steps, steps_ref), rel_l2(steps[big], steps_ref[big])
    print(f"fused Adam first-step update vs torch.optim.Adam: rel-L2 {e_all:.2e} over all {steps.size} parameters, "
          f"{e_big:.2e} over the {int(big.sum())} with |grad| > 1e-6")
    assert e_all < 5e-2 and e_big < 1e-2
    # the sampler sees the updated weights
    y = m.denoise(data.cuda(), E=E.cuda(), sigma=torch.ones(2).cuda(), layers=layers.cuda())
    assert torch.isfinite(y).all()


def test_dropin_training_loop_decreases_loss():
    """train/train_diffusion.py:42-66 verbatim shape: zero_grad -> compute_loss -> backward -> torch optimizer step."""
    m, c = _model("ds2", max_batch=4)
    opt = torch.optim.Adam(m.parameters(), lr=2e-3)
    g = torch.Generator().manual_seed(3)
    data = torch.randn([4] + list(c["shape"]), generator=g).cuda()
    E = torch.rand(4, 1, generator=g).cuda()
    layers = torch.randn(4, 46, generator=g).cuda()
    noise = torch.randn([4] + list(c["shape"]), generator=g).cuda()
    rnd = torch.randn(4, generator=g).cuda()
    losses = []
    for _ in range(6):
        m.zero_grad()
        opt.zero_grad()
        loss = m.compute_loss(data, E, noise=noise, layers=layers, rnd_normal=rnd)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    print("drop-in loop losses:", [f"{v:.4e}" for v in losses])
    assert losses[-1] < losses[0] and np.isfinite(losses).all()


def test_grad_sync_hook_and_trainer_single_process():
    """The data-parallel hooks on one GPU: `grad_sync` sees the flat gradient right after the backward pass and its return
    value divides the gradient (both in loss.backward() and in the fused Adam); DataParallelTrainer.step without a process
    group equals compute_loss + fused_adam_step; read_loss=False returns None and leaves the same gradient."""
    from calodiffusion_b200.distributed import DataParallelTrainer
    m, c = _model("ds2")
    E, layers, data, noise, rnd, time = [None if t is None else t.cuda() for t in train_inputs(c["shape"], 2, 1, 46)]
    loss = m.loss_function(m, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    loss.backward()
    g1 = torch.cat([p.grad.reshape(-1) for p in m.model.parameters()]).clone()
    G1 = m.grad_buffer().clone()
    seen = []

    def fake_sync(flat):
        seen.append(flat.data_ptr())
        flat.mul_(3.0)              # "sum over 3 identical ranks"
        return 3.0
    m.grad_sync = fake_sync
    for p in m.model.parameters():
        p.grad = None
    loss2 = m.loss_function(m, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    loss2.backward()
    g2 = torch.cat([p.grad.reshape(-1) for p in m.model.parameters()])
    assert seen == [m.grad_buffer().data_ptr()]
    e = float((g2 - g1).norm() / g1.norm())
    print(f"gradient through a x3 / 3 grad_sync vs plain: rel-L2 {e:.2e}")
    assert e < 2e-3                  # run-to-run noise of the fp16 / atomic accumulation orders (measured 3.6e-4), not the x3 / 3
    assert m.loss_function(m, data, E, noise=noise, layers=layers, rnd_normal=rnd) is not None
    m._read_loss = False
    assert m.loss_function(m, data, E, noise=noise, layers=layers, rnd_normal=rnd) is None
    m._read_loss = True
    assert float((m.grad_buffer() / 3.0 - G1).norm() / G1.norm()) < 2e-3
    # trainer without a process group: no collective, plain fused Adam
    m2, _ = _model("ds2")
    m3, _ = _model("ds2")
    tr = DataParallelTrainer(m2, lr=1e-3)
    l2 = tr.step(data, E, layers, noise=noise, rnd_normal=rnd)
    l3 = m3.loss_function(m3, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    m3.fused_adam_step(1e-3)
    assert abs(float(l2) - float(l3)) < 1e-4 * abs(float(l3))      # run-to-run (atomic orders): measured 1.7e-5
    diff = (m2._train["P"] - m3._train["P"]).abs()
    frac = float((diff > 1e-5).float().mean())
    print(f"trainer vs manual step: parameters differing by > 1e-5: {frac:.2e} of all (max {float(diff.max()):.2e})")
    assert frac < 0.02                # Adam's first step is +-lr per element: a sign can flip only where a gradient is ~0


def test_training_batch_smaller_than_max_batch_and_sampling_after_training():
    """(1) A training handle sized for max_batch = 4 stepping on B = 2 gives the gradient of the max_batch = 2 handle (every
    per-sample scratch offset scales with the CURRENT batch).  (2) After a fused Adam step the sampling engine sees the updated
    master parameters: `denoise` equals a fresh model loaded from `state_dict()`."""
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    m2, c = _model("ds2", max_batch=2)
    m4, _ = _model("ds2", max_batch=4)
    E, layers, data, noise, rnd, time = [None if t is None else t.cuda() for t in train_inputs(c["shape"], 2, 1, 46)]
    l2 = m2.loss_function(m2, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    l4 = m4.loss_function(m4, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    assert abs(float(l2) - float(l4)) < 1e-4 * abs(float(l2))
    g2, g4 = m2.grad_buffer(), m4.grad_buffer()
    e = float((g2 - g4).norm() / g2.norm())
    print(f"B = 2 on a max_batch = 4 handle vs a max_batch = 2 handle: gradient rel-L2 {e:.2e}")
    assert e < 2e-3                  # run-to-run noise of the atomic accumulation orders (measured 3.5e-4)
    # (2)
    before = m4.denoise(data, E=E, sigma=torch.tensor([1.5, 0.3]).cuda(), layers=layers).clone()
    m4.loss_function(m4, data, E, noise=noise, layers=layers, rnd_normal=rnd)
    m4.fused_adam_step(1e-2)
    after = m4.denoise(data, E=E, sigma=torch.tensor([1.5, 0.3]).cuda(), layers=layers)
    cfg = dict(c["cfg"], SAMPLER="DDPM")
    fresh = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=2)
    fresh.load_state_dict({k: v.clone() for k, v in m4.state_dict().items()})
    ref = fresh.denoise(data, E=E, sigma=torch.tensor([1.5, 0.3]).cuda(), layers=layers)
    moved = rel_l2(after.cpu().numpy(), before.cpu().numpy())
    same = rel_l2(after.cpu().numpy(), ref.cpu().numpy())
    print(f"denoise after one fused Adam step: moved {moved:.2e} from the old weights, {same:.2e} from a fresh model with state_dict()")
    assert moved > 1e-3 and same < 1e-4           # measured 7.7e-6 (weights packed on the device vs on the host)
