"""GPU parity of FerroelectricBasisConv2d / MemEfficient_FerroelectricBasisConv2d (csrc/ferro.cu, implicit im2col)
through the C ABI: golden vectors from the live reference classes (oracle/make_golden_conv.py), the oracle on a larger
shape, and gradients (image, every parameter, out_bias) against fp64 oracle autograd."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from oracle import layers as OL

pytestmark = pytest.mark.gpu
TOL, GRAD_TOL = 1e-5, 2e-4

CASES = (("plain_s1p1", fetode.FerroelectricBasisConv2d, 5, dict(kernel_size=3, stride=1, padding=1)),
         ("plain_s2p0_c1", fetode.FerroelectricBasisConv2d, 1, dict(kernel_size=(3, 2), stride=2, padding=0)),
         ("memeff_s1p1", fetode.MemEfficient_FerroelectricBasisConv2d, 5, dict(kernel_size=3, stride=1, padding=1, out_chunk=2)))


@pytest.mark.parametrize("tag,cls,cout,kw", CASES)
def test_conv2d_golden_stateless_and_stateful(cuda, tag, cls, cout, kw):
    p, d = load_golden("ferroconv_" + tag)
    m = cls(2, cout, num_basis=3, **kw)
    missing, unexpected = m.load_state_dict(p, strict=False)
    assert not unexpected and all(k.startswith("_") for k in missing)
    m = m.to(cuda)
    with torch.no_grad():
        for i in (0, 1):
            assert max_rel(m(d[f"x{i}"].to(cuda)), d[f"stateless{i}"]) < TOL
        m.stateful = True
        m.reset_state()
        if "stateful0" in d:
            for i in range(3):
                assert max_rel(m(d[f"x{i}"].to(cuda)), d[f"stateful{i}"]) < TOL, i
        else:
            with pytest.raises(RuntimeError):
                m(d["x0"].to(cuda))


@pytest.mark.parametrize("mem_eff,stateful,cout", [(False, False, 12), (True, True, 12), (False, True, 1)])
def test_conv2d_values_and_gradients_vs_oracle(cuda, mem_eff, stateful, cout):
    gen = torch.Generator().manual_seed(3 + cout)
    cls = fetode.MemEfficient_FerroelectricBasisConv2d if mem_eff else fetode.FerroelectricBasisConv2d
    torch.manual_seed(1)
    m = cls(4, cout, kernel_size=3, num_basis=3, stride=2, padding=1, stateful=stateful)
    with torch.no_grad():
        m.out_bias.copy_(torch.randn(cout, generator=gen))
    p = {k: v.detach().clone() for k, v in m.state_dict().items() if not k.startswith("_")}
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    m = m.to(cuda)
    xs = [torch.randn(5, 4, 11, 9, generator=gen) * 1.5 for _ in range(2)]
    st64, st32 = OL.FerroConvState(), OL.FerroConvState()
    kw = dict(stride=2, padding=1, stateful=stateful, mem_efficient=mem_eff)
    x64 = [x.double().requires_grad_(True) for x in xs]
    outs64 = [OL.ferro_conv2d_forward(x, p64, st64, **kw) for x in x64]
    ws = [torch.randn(o.shape, generator=gen, dtype=torch.float64) for o in outs64]
    sum((o * w).sum() for o, w in zip(outs64, ws)).backward()
    with torch.no_grad():
        want32 = [OL.ferro_conv2d_forward(x, p, st32, **kw) for x in xs]
    xg = [x.to(cuda).requires_grad_(True) for x in xs]
    outs = [m(x) for x in xg]
    for o, w32 in zip(outs, want32):
        assert o.shape == w32.shape and max_rel(o, w32) < TOL
    sum((o * w.float().to(cuda)).sum() for o, w in zip(outs, ws)).backward()
    for a, b in zip(xg, x64):
        assert max_rel(a.grad, b.grad) < GRAD_TOL
    for name, t in m.named_parameters():
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name


def test_conv2d_return_activations_layout(cuda):
    p, d = load_golden("ferroconv_plain_s1p1")
    m = fetode.FerroelectricBasisConv2d(2, 5, 3, num_basis=3, stride=1, padding=1)
    m.load_state_dict(p, strict=False)
    m = m.to(cuda)
    with torch.no_grad():
        out, basis, coef = m(d["x0"].to(cuda), return_activations=True)
        _, want = OL.ferro_conv2d_forward(d["x0"], p, stride=1, padding=1, return_basis=True)
    assert basis.shape == want.shape and max_rel(basis, want) < TOL
    assert coef.shape == (1, 5, 2, 3, 3, 3, 1)
