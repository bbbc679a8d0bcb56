import os, sys, json
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, fetode
dev = torch.device("cuda:0")
for (B, I, O) in ((4096, 768, 128), (4096, 128, 128), (4096, 2352, 128)):
    torch.manual_seed(0)
    layer = fetode.KANLinear(I, O, num_basis=10).to(dev)
    x = torch.randn(B, I, device=dev, requires_grad=True)
    g = torch.randn(B, O, device=dev)
    out = {"shape": (B, I, O)}
    for mode in ("0", "1"):
        os.environ["FETODE_TC"] = mode
        def step():
            layer.zero_grad(); x.grad = None
            y = layer(x); y.backward(g)
        for _ in range(3): step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); [step() for _ in range(5)]; e1.record(); torch.cuda.synchronize()
        out["fwd_bwd_ms_tc" + mode] = e0.elapsed_time(e1) / 5
    print(json.dumps(out), flush=True)
