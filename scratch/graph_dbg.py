import sys, os, warnings
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "fet-ode_b200"))
import torch, torch.nn.functional as F, fetode
from fetode import ecg
warnings.simplefilter("error")
dev = torch.device("cuda")
torch.manual_seed(0)
m = ecg.KanFet_MLP_NODE(T=24, num_classes=2, latent_dim=16, num_basis=6, ode_hidden=32, dropout=0.0)
m.ode_method = "rk4_classic"
m = m.to(dev)
x = torch.randn(48, 24, device=dev); y = torch.randint(0, 2, (48,), device=dev)
opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
for _ in range(2):
    opt.zero_grad(set_to_none=True); F.cross_entropy(m(x), y).backward(); opt.step()
print("eager ok")
gopt = fetode.optim.AdamW(m.parameters(), lr=1e-3, capturable=True)
g = fetode.graphs.GraphedTrainStep(m, gopt, F.cross_entropy, x, y, max_norm=1.0, warmup=2)
print("graph ok", float(g(x, y)))
