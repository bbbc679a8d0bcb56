import os, sys
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
os.environ["FETODE_TC"] = "1"
import torch, fetode
dev = torch.device("cuda:0")
torch.manual_seed(0)
I, O, B = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
layer = fetode.KANLinear(I, O, num_basis=10).to(dev)
x = torch.randn(B, I, device=dev, requires_grad=True)
g = torch.randn(B, O, device=dev)
y = layer(x); y.backward(g)
torch.cuda.synchronize()
gx1 = x.grad.clone()
os.environ["FETODE_TC"] = "0"
layer.zero_grad(); x.grad = None
y = layer(x); y.backward(g)
torch.cuda.synchronize()
print("ok", I, O, B, float((gx1 - x.grad).abs().max() / x.grad.abs().max()))
