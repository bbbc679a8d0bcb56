"""LV training iteration with torchdiffeq's default METHOD and tolerances (dopri5, rtol 1e-7 / atol 1e-9: the call of
train_kanfet_node_predprey.py:252) -- fused attempt kernels vs the staged path, fwd+bwd+Adam, wall-clock per iteration
(the adaptive loop reads the controller state back, so the step is not graph-capturable).
usage: python profiles/experiments/dopri5_time.py [B ...]   (env RTOL / ATOL / BRANCH=KAN|KANFET / ONLY=fused /
FETODE_DOPRI5_DEVICE=0 for host-controlled attempt launches)"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "fet-ode_b200"))
import fetode  # noqa: E402


def run(B, fused, iters=5):
    os.environ["FETODE_DOPRI5_FUSED"] = "1" if fused else "0"
    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    m = getattr(fetode, os.environ.get("BRANCH", "KANFET"))([2, 10, 2]).to(dev)
    rtol, atol = float(os.environ.get("RTOL", "1e-7")), float(os.environ.get("ATOL", "1e-9"))
    t = torch.linspace(0, 3.5, 35, dtype=torch.float64)
    x0 = torch.ones(B, 2, device=dev) if B == 1 else (torch.rand(B, 2, device=dev) * 6.75 + 0.25)
    tgt = torch.randn(35, B, 2, device=dev)
    opt = fetode.optim.Adam(m.parameters(), lr=2e-3)
    st = {}

    def it():
        opt.zero_grad()
        if hasattr(m, "reset_state"):
            m.reset_state()
        pred = fetode.odeint(lambda tt, X: m(X), x0, t, rtol=rtol, atol=atol, options={"stats": st})
        ((pred - tgt) ** 2).mean().backward()
        opt.step()
    for _ in range(2):
        it()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        it()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / iters * 1e3
    return {"B": B, "device_controller": bool(st.get("device_controller")), "model": type(m).__name__, "rtol": rtol, "atol": atol, "fused": fused, "ms_per_iteration": ms, "accepted": st.get("accepted"), "rejected": st.get("rejected")}


if __name__ == "__main__":
    for B in [int(a) for a in sys.argv[1:]] or [1, 4096]:
        for fused in ((True,) if os.environ.get("ONLY") == "fused" else (True, False)):
            print(json.dumps(run(B, fused, iters=5 if fused else 2)), flush=True)
