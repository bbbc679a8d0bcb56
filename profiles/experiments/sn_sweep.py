#!/usr/bin/env python
"""ms per launch of the fused LV KANFET([2,10,2]) Runge-Kutta kernel.  usage: sn_sweep.py B n_steps [branch]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import fetode  # noqa: E402
from fetode._lib import TABLEAU  # noqa: E402

B, n = int(sys.argv[1]), int(sys.argv[2])
branch = sys.argv[3] if len(sys.argv) > 3 else "ferro"
torch.manual_seed(0)
m = (fetode.KANFET if branch == "ferro" else fetode.KAN)([2, 10, 2], grid_size=5).cuda()
plan = m.small_plan()
gen = torch.Generator().manual_seed(1)
x0 = (torch.rand(B, 2, generator=gen) * 6.75 + 0.25).cuda()
y = torch.empty_like(x0)
st = torch.empty(B, max(plan.state_floats, 1), device="cuda")
plan.image()
run = lambda: plan.rk_fwd(TABLEAU["rk4_classic"], x0, 0.014, n, state=st, state_is_zero=True, y_out=y)
for _ in range(2):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
it = 3
e0.record()
for _ in range(it):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / it
print(json.dumps({"B": B, "steps": n, "branch": branch, "shape": os.environ.get("FETODE_SN2_SHAPE", "auto"),
                  "v1": os.environ.get("FETODE_SN_V1", "0"), "ms": ms, "traj_steps_per_s": B * n / ms * 1e3,
                  "finite": bool(torch.isfinite(y).all()), "checksum": float(y.double().sum())}))
