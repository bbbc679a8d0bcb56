import os, sys, json
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, torch.nn.functional as F
import fetode
from fetode import ecg
from torch.profiler import profile, ProfilerActivity
DEV = torch.device("cuda:0")
which = sys.argv[1]
if which == "ecg":
    torch.manual_seed(0)
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    model = model.to(DEV)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    B = 4096
    x = torch.randn(B, 96, device=DEV); y = torch.randint(0, 2, (B,), device=DEV)
    def step():
        opt.zero_grad(set_to_none=True)
        F.cross_entropy(model(x), y).backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
elif which == "lv":
    torch.manual_seed(0)
    model = fetode.KANFET([2, 10, 2], grid_size=5).to(DEV)
    opt = torch.optim.Adam(model.parameters(), lr=2e-3)
    B = int(sys.argv[2]); n = 34
    t = torch.linspace(0, 3.5, n + 1, dtype=torch.float64)
    x0 = (torch.rand(B, 2) * 6.75 + 0.25).to(DEV); tgt = torch.randn(n + 1, B, 2, device=DEV)
    def step():
        model.reset_state()
        opt.zero_grad(set_to_none=True)
        pred = fetode.odeint(lambda tt, X: model(X), x0, t, method="rk4_classic")
        ((pred - tgt) ** 2).mean().backward()
        opt.step()
elif which == "ett":
    torch.manual_seed(0)
    enc = torch.nn.Sequential(torch.nn.Flatten(), fetode.KAN([336 * 7, 128]), torch.nn.ReLU(), fetode.KAN([128, 64])).to(DEV)
    x = torch.randn(4096, 336, 7, device=DEV)
    def step():
        enc.zero_grad(set_to_none=True)
        enc(x).square().mean().backward()
for _ in range(2): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=70))
