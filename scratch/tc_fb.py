import os, sys
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
os.environ["FETODE_TC"] = "1"
import torch, fetode
dev = torch.device("cuda:0")
torch.manual_seed(0)
layer = fetode.KANLinear(768, 128, num_basis=10).to(dev)
x = torch.randn(4096, 768, device=dev, requires_grad=True)
g = torch.randn(4096, 128, device=dev)
for _ in range(3):
    layer.zero_grad(); x.grad = None
    y = layer(x); y.backward(g)
torch.cuda.synchronize()
print("ok", float(y.abs().max()), float(x.grad.abs().max()))
