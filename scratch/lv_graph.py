import sys, os, copy
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "fet-ode_b200"))
import torch, fetode
dev = torch.device("cuda")
for B in (1, 65536):
    torch.manual_seed(0)
    m1 = fetode.KANFET([2, 10, 2], grid_size=5)
    m2 = copy.deepcopy(m1)
    m1, m2 = m1.to(dev), m2.to(dev)
    n, T = 34, 3.5
    t = torch.linspace(0, T, n + 1, dtype=torch.float64)
    gen = torch.Generator().manual_seed(1)
    x0 = ((torch.rand(B, 2, generator=gen) * 6.75 + 0.25) if B > 1 else torch.ones(1, 2)).to(dev)
    tgt = torch.randn(n + 1, B, 2, generator=gen).to(dev)
    def fwd(m, x):
        m.reset_state()
        return fetode.odeint(lambda tt, X: m(X), x, t, method="rk4_classic")
    mse = lambda pred, y: ((pred - y) ** 2).mean()
    o1 = fetode.optim.Adam(m1.parameters(), lr=2e-3)
    p0 = [p.detach().clone() for p in m1.parameters()]
    L1 = []
    for _ in range(6):
        o1.zero_grad(); l1 = mse(fwd(m1, x0), tgt); l1.backward(); o1.step(); L1.append(float(l1.detach()))
    o2 = fetode.optim.Adam(m2.parameters(), lr=2e-3, capturable=True)
    g = fetode.graphs.GraphedTrainStep(m2, o2, mse, x0, tgt, warmup=3, forward=fwd)
    L2 = []
    for _ in range(3):
        l2 = g(x0, tgt); L2.append(float(l2))
    print("eager losses", L1, "graph losses", L2, "moved", max(float((p - q).abs().max()) for p, q in zip(m1.parameters(), p0)))
    err = max(float((p - q).abs().max() / (p.abs().max() + 1e-12)) for p, q in zip(m1.parameters(), m2.parameters()))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        g(g.x, g.y)
    e1.record(); torch.cuda.synchronize()
    print(B, "loss eager/graph", float(l1.detach()), float(l2), "param rel err", err, "graph ms", e0.elapsed_time(e1) / 20)
