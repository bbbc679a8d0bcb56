#!/usr/bin/env python
"""us per call of the dense head products (fetode_linear_fwd / _bwd, csrc/linear.cu) at the ECG shape [4096,128] x [128,64],
back-to-back launches on one stream (warm caches, as inside the replayed training step).
usage: linear_time.py [path/to/libfetode.so]"""
import ctypes as C
import json
import sys

import torch

import os
lib = C.CDLL(os.path.abspath(sys.argv[1] if len(sys.argv) > 1 else "fet-ode_b200/fetode/libfetode.so"))
dev = torch.device("cuda:0")
B, K, O = 4096, 128, 64
x = torch.randn(B, K, device=dev); w = torch.randn(O, K, device=dev); b = torch.randn(O, device=dev)
y = torch.empty(B, O, device=dev); gy = torch.randn(B, O, device=dev)
gx = torch.empty(B, K, device=dev); gw = torch.zeros(O, K, device=dev); gb = torch.zeros(O, device=dev)
P = lambda t: C.c_void_p(t.data_ptr())
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
fwd = lambda: lib.fetode_linear_fwd(P(x), P(w), P(b), P(y), C.c_int64(B), C.c_int32(K), C.c_int32(O), st)
bwd = lambda: lib.fetode_linear_bwd(P(x), P(w), P(gy), P(gx), P(gw), P(gb), C.c_int64(B), C.c_int32(K), C.c_int32(O), st)
out = {"lib": sys.argv[1] if len(sys.argv) > 1 else "in-tree"}
for name, fn in (("fwd", fwd), ("bwd", bwd)):
    for _ in range(20):
        assert fn() == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 500
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    out[name + "_us"] = e0.elapsed_time(e1) / n * 1e3
ref = torch.nn.functional.linear(x.double(), w.double(), b.double())
out["fwd_max_rel_vs_fp64"] = float((y - ref).abs().max() / ref.abs().max())
gw.zero_(); gb.zero_(); bwd(); torch.cuda.synchronize()
out["gx_max_rel_vs_fp64"] = float((gx - gy.double() @ w.double()).abs().max() / (gy.double() @ w.double()).abs().max())
out["gw_max_rel_vs_fp64"] = float((gw - gy.double().t() @ x.double()).abs().max() / (gy.double().t() @ x.double()).abs().max())
print(json.dumps(out))
