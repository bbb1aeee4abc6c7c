"""cd_conv_noalloc (include/calodiffusion_b200.h): the allocation-free convolution entry point against cd_test_conv's SIMT
backend on the same inputs, through the C-ABI, twice in a row (the static scratch array keeps the first call's contents).
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import load_case, rel_l2

pytestmark = pytest.mark.gpu

TOL_LAYER = 2e-3      # tests/test_gpu_parity.py: one layer vs torch fp32
TOL_SAME = 1e-6       # same kernel, same fp16 staging, same fp32 weights: identical up to nothing at all (measured 0)


def _model():
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case("ds2")
    cfg = dict(c["cfg"], SAMPLER="DDPM")
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=2)
    m.load_state_dict(dict(c["sd"]))
    return m


def _shapes(lib, h, li, lo):
    zi, zo = (C.c_int * 3)(), (C.c_int * 3)()
    lib.cd_level_shape(h, li, zi)
    lib.cd_level_shape(h, lo, zo)
    return list(zi), list(zo)


def _both(m, lib, li, lo, Cin, Cout, k, stride, transposed, seed, B=2):
    h = m._engine(B)
    zi, zo = _shapes(lib, h, li, lo)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Cin, *zi, generator=g).half().float()
    wshape = (Cin, Cout, *k) if transposed else (Cout, Cin, *k)
    w = (torch.randn(wshape, generator=g) / np.sqrt(Cin * np.prod(k))).half().float().contiguous()
    b = torch.randn(Cout, generator=g).contiguous()
    xd, wd, bd = x.cuda().contiguous(), w.cuda(), b.cuda()
    ks, st = (C.c_int * 3)(*k), (C.c_int * 3)(*stride)
    y_old = torch.zeros(B, Cout, *zo, device="cuda")
    rc = lib.cd_test_conv(h, li, lo, Cin, Cout, ks, st, int(transposed), w.data_ptr(), b.data_ptr(), xd.data_ptr(), y_old.data_ptr(),
                          B, 1, None)
    assert rc == 0, lib.cd_last_error().decode()
    y_new = torch.full((B, Cout, *zo), float("nan"), device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free0 = torch.cuda.mem_get_info()[0]
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.cd_conv_noalloc(h, li, lo, Cin, Cout, ks, st, int(transposed), wd.data_ptr(), bd.data_ptr(), xd.data_ptr(),
                             y_new.data_ptr(), B, stream)
    assert rc == 0, lib.cd_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return y_old.cpu().numpy(), y_new.cpu().numpy(), free0 - torch.cuda.mem_get_info()[0]


def test_no_alloc_conv_matches_test_conv_twice_in_a_row():
    m = _model()
    lib = m._lib
    assert lib.cd_conv_noalloc_max_bytes() == 16 << 20
    # two calls back to back with DIFFERENT shapes, so the second one runs over the first one's leftovers in the shared array;
    # then the first shape again
    cases = [
        (0, 0, 32, 32, (3, 3, 3), (1, 1, 1), False, 1),
        (0, 1, 32, 16, (3, 4, 4), (2, 2, 2), False, 2),
        (0, 0, 32, 32, (3, 3, 3), (1, 1, 1), False, 3),
    ]
    for n, (li, lo, Cin, Cout, k, st, tr, seed) in enumerate(cases):
        y_old, y_new, used = _both(m, lib, li, lo, Cin, Cout, k, st, tr, seed)
        if n:       # (the first call may page in the kernels' code: lazy module loading)
            assert used == 0, f"device memory in use grew by {used} bytes across cd_conv_noalloc"
        assert np.isfinite(y_new).all()
        e = rel_l2(y_new, y_old)
        print(f"no-alloc conv L{li}->L{lo} {Cin}->{Cout} k{k}: rel-L2 vs cd_test_conv = {e:.2e}")
        assert e < TOL_SAME, e
        assert e < TOL_LAYER


def test_no_alloc_conv_rejects_oversize_and_bad_arguments():
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    m = _model()
    lib = m._lib
    h = m._engine(2)
    zi, zo = _shapes(lib, h, 0, 0)
    ks, st = (C.c_int * 3)(3, 3, 3), (C.c_int * 3)(1, 1, 1)
    x = torch.zeros(2, 8, *zi, device="cuda")
    y = torch.zeros(2, 8, *zo, device="cuda")
    w = torch.zeros(8, 8, 3, 3, 3, device="cuda")
    assert lib.cd_conv_noalloc(h, 0, 0, 12, 8, ks, st, 0, w.data_ptr(), None, x.data_ptr(), y.data_ptr(), 2, None) != 0
    assert "multiples of 8" in lib.cd_last_error().decode()
    assert lib.cd_conv_noalloc(h, 0, 0, 8, 8, ks, st, 0, None, None, x.data_ptr(), y.data_ptr(), 2, None) != 0
    # beyond the static array: the staging tensors scale with the handle's max_batch; nothing is launched, nothing dereferenced
    c = load_case("ds2")
    cfg = dict(c["cfg"], SAMPLER="DDPM")
    big = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=64)
    big.load_state_dict(dict(c["sd"]))
    hb = big._engine(64)
    rc = lib.cd_conv_noalloc(hb, 0, 0, 64, 64, ks, st, 0, w.data_ptr(), None, x.data_ptr(), y.data_ptr(), 2, None)
    assert rc != 0
    msg = lib.cd_last_error().decode()
    print(msg)
    assert "scratch" in msg and "16777216" in msg
    torch.cuda.synchronize()
