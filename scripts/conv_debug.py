"""Timing breakdown of one R-stacked conv launch (needs a library built with -DCD_MMA_DEBUG)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import WORKLOADS
from calodiffusion_b200.calodiffusion import CaloDiffusion
wl = WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "ds2"]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 600
Cin, Cout = int(sys.argv[3]) if len(sys.argv) > 3 else 32, int(sys.argv[4]) if len(sys.argv) > 4 else 32
cfg = dict(wl["cfg"])
m = CaloDiffusion(cfg, n_steps=400, loss_type="l2", max_batch=B)
h = m._engine(B)
lib = m._lib
z = (C.c_int * 3)(); lib.cd_level_shape(h, 0, z)
x = torch.randn(B, Cin, *list(z), device="cuda")
y = torch.zeros(B, Cout, *list(z), device="cuda")
w = (torch.randn(Cout, Cin, 3, 3, 3) / 30).contiguous(); b = torch.randn(Cout).contiguous()
ks = (C.c_int * 3)(3, 3, 3); st = (C.c_int * 3)(1, 1, 1)
out = (C.c_longlong * 16)()
for mode in ([int(a) for a in sys.argv[5].split(',')] if len(sys.argv) > 5 else (0, 2)):
    os.environ["CD_DBG_MODE"] = str(mode)
    for rep in range(2):
        lib.cd_debug_counters(out, 1)
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        rc = lib.cd_test_conv(h, 0, 0, Cin, Cout, ks, st, 0, w.data_ptr(), b.data_ptr(), x.data_ptr(), y.data_ptr(), B, 0, None)
        assert rc == 0, lib.cd_last_error()
        ev1.record()
        torch.cuda.synchronize()
    print(f"mode {mode}: whole test call (conversions + conv) {ev0.elapsed_time(ev1):.3f} ms")
    lib.cd_debug_counters(out, 0)
    v = list(out)
    sub = max(v[6], 1) * 2
    print(f"mode {mode}: issuer wait a_full {v[0]/max(v[3],1):.0f}/tile  wait t_empty {v[1]/sub:.0f}/sub  issue {v[2]/sub:.0f}/sub | "
          f"epi wait {v[4]/max(v[6],1):.0f}  work {v[5]/max(v[6],1):.0f} per own subtile (cumulative: tmem ld {v[9]/max(v[6],1):.0f}, bar {v[10]/max(v[6],1):.0f}, shuffles {v[11]/max(v[6],1):.0f}, stores {v[12]/max(v[6],1):.0f}); tiles {v[3]} subtiles {sub}")
