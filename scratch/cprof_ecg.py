import os, sys, cProfile, pstats, io
sys.path[:0] = ["/root/repo", "/root/repo/fet-ode_b200"]
import torch, torch.nn.functional as F
from fetode import ecg
DEV = torch.device("cuda:0")
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
for _ in range(3): step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(3): step()
torch.cuda.synchronize()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(45)
print(s.getvalue())
