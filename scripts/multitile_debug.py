import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from test_gpu_parity import _model, _conv_case
m, c = _model("ds2", max_batch=48)
lib = m._lib
cases = {"k3_l0_32_32": (0,0,32,32,(3,3,3),24), "k1_l0_32_96": (0,0,32,96,(1,1,1),24), "k3_l0_64_32_kstaged": (0,0,64,32,(3,3,3),24),
         "k3_l1_64_64_stream": (1,1,64,64,(3,3,3),48), "k3_l1_128_32_B8": (1,1,128,32,(3,3,3),8), "k3_l1_128_32_B20": (1,1,128,32,(3,3,3),20), "k3_l1_128_32_B48": (1,1,128,32,(3,3,3),48)}
for k,(li,lo,ci,co,ks,B) in cases.items():
    try:
        e = _conv_case(m, lib, li, lo, ci, co, ks, (1,1,1), False, 0, B=B)
        print(k, f"{e:.2e}", flush=True)
    except AssertionError as ex:
        print(k, "FAILED", str(ex)[:100], flush=True); break
