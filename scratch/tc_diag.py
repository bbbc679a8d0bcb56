import os, sys, json
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, fetode
from oracle import layers as OL
dev = torch.device("cuda:0")
gen = torch.Generator().manual_seed(3)
I, O, nb, B = 768, 128, 10, 4096
p = OL.make_kanlinear_params(I, O, gen, nb=nb, branch="logistic")
layer = fetode.KANLinear(I, O, num_basis=nb); layer.load_state_dict(p, strict=False); layer = layer.to(dev)
x = (torch.randn(B, I, generator=gen) * 0.9)
p64 = {k: v.double().to(dev) for k, v in p.items()}
with torch.no_grad():
    want = OL.kanlinear_forward(x.double().to(dev), p64, branch="logistic")
    p32 = {k: v.to(dev) for k, v in p.items()}
    ref32 = OL.kanlinear_forward(x.to(dev), p32, branch="logistic")
    def rel(a): return float((a.double() - want).abs().max() / want.abs().max())
    print("torch fp32 (cuda) vs fp64:", rel(ref32))
    xd = x.to(dev)
    os.environ["FETODE_TC"] = "0"; print("cuda-core vs fp64:", rel(layer(xd)))
    os.environ["FETODE_TC"] = "1"
    for ups in (1, 2, 3, 6, 12, 24):
        os.environ["FETODE_TC_UPS"] = str(ups)
        y = layer(xd); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); [layer(xd) for _ in range(5)]; e1.record(); torch.cuda.synchronize()
        err = (y.double() - want)
        print(f"tc ups={ups}: rel {rel(y):.3e} mean signed err {float(err.mean()):.3e} mean|err| {float(err.abs().mean()):.3e}  ms {e0.elapsed_time(e1)/5:.3f}")
