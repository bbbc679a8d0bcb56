import os, sys, ctypes as C
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, fetode
from fetode import ops
from fetode._lib import lib, ptr, stream, KanLayerGrads, check
dev = torch.device("cuda:0")
torch.manual_seed(0)
I, O, B = 64, 128, 128
layer = fetode.KANLinear(I, O, num_basis=10).to(dev)
x = torch.randn(B, I, device=dev)
g = torch.randn(B, O, device=dev)
s, blob = layer._pack.get()
L = lib()
n_ws = L.fetode_kanlinear_bwd_ws_floats(C.byref(s))
out = {}
for mode in ("0", "1"):
    os.environ["FETODE_TC"] = mode
    ws = torch.full((n_ws,), float("nan"), device=dev)
    gr = KanLayerGrads()
    gx = torch.empty_like(x)
    check(L.fetode_kanlinear_bwd(C.byref(s), ptr(blob), ptr(x), None, C.c_int(1), ptr(g), ptr(gx), C.byref(gr), ptr(ws), C.c_int64(B), stream()))
    torch.cuda.synchronize()
    out[mode] = ws[: I * 19 * O].view(I, 19, O).clone()
a, b = out["1"], out["0"]
print("ref absmax", float(b.abs().max()), "tc absmax", float(a.abs().max()), "nan tc", int(torch.isnan(a).sum()))
print("maxdiff", float((a - b).abs().max()))
print("tc  [0,:,0]", a[0, :, 0])
print("ref [0,:,0]", b[0, :, 0])
print("tc  [5,:,77]", a[5, :, 77])
print("ref [5,:,77]", b[5, :, 77])
