import os, sys
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, fetode
dev = torch.device("cuda:0")
torch.manual_seed(0)
I, O, B = 64, 128, 128
layer = fetode.KANLinear(I, O, num_basis=10).to(dev)
x = torch.randn(B, I, device=dev)
g = torch.randn(B, O, device=dev)
res = {}
for mode in ("0", "1"):
    os.environ["FETODE_TC"] = mode
    layer.zero_grad()
    xs = x.clone().requires_grad_(True)
    (layer(xs) * g).sum().backward()
    res[mode] = {n: t.grad.clone() for n, t in layer.named_parameters()}
for n in res["0"]:
    a, b = res["1"][n], res["0"][n]
    print(n, tuple(a.shape), "tc absmax %.4g ref absmax %.4g  maxdiff %.4g  nonzero frac %.3f" % (a.abs().max(), b.abs().max(), (a - b).abs().max(), (a != 0).float().mean()))
a, b = res["1"]["base_weight"], res["0"]["base_weight"]
print("base_weight tc[:4,:6]\n", a[:4, :6], "\nref\n", b[:4, :6])
sw = res["1"]["spline_weight"]; print("spline tc nonzero per input (first 16):", (sw != 0).float().mean(dim=(0, 2))[:16])
