#!/usr/bin/env python
"""One training step of a named configuration, for ncu launch lists: step_once.py lv|ecg|mnist [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import fetode  # noqa: E402
from fetode import ecg, mnist_kuramoto_kan as MK  # noqa: E402

which = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
if which == "lv":
    model = fetode.KANFET([2, 10, 2], grid_size=5).to(dev)
    opt = fetode.optim.Adam(model.parameters(), lr=2e-3)
    B, n = int(os.environ.get("LV_ROWS", "65536")), 34
    t = torch.linspace(0, 3.5, n + 1, dtype=torch.float64)
    x0 = (torch.rand(B, 2) * 6.75 + 0.25).to(dev)
    tgt = torch.randn(n + 1, B, 2).to(dev)
    func = fetode.as_rhs(model)

    def step():
        model.reset_state()
        opt.zero_grad()
        ((fetode.odeint(func, x0, t, method="rk4_classic") - tgt) ** 2).mean().backward()
        opt.step()
elif which == "ecg":
    model = ecg.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128, dropout=0.0)
    model.ode_method = "rk4_classic"
    model = model.to(dev)
    opt = fetode.optim.AdamW(model.parameters(), lr=1e-3)
    x = torch.randn(4096, 96, device=dev)
    y = torch.randint(0, 2, (4096,), device=dev)

    def step():
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        opt.clip_grad_norm_(1.0)
        opt.step()
else:
    model = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8).to(dev)
    opt = fetode.optim.AdamW(model.parameters(), lr=1e-3)
    x = torch.rand(1024, 1, 28, 28, device=dev)
    y = torch.randint(0, 10, (1024,), device=dev)

    def step():
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        opt.step()
for _ in range(iters):
    step()
torch.cuda.synchronize()
print("ok")
