"""GPU parity of the Kuramoto lattice kernels (csrc/kuramoto.cu) and the MNIST classifier head, through the C ABI."""
import pytest
import torch
import torch.nn.functional as F

from conftest import load_golden, max_rel
from fetode import mnist_kuramoto_kan as MK
from oracle import layers as OL

pytestmark = pytest.mark.gpu
RHS_TOL = 1e-5
GRAD_TOL = 2e-4


def _clf(params, cuda):
    m = MK.KuramotoKANClassifier(kuramoto_steps=10, num_basis=8)
    missing, unexpected = m.load_state_dict(params, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return m.to(cuda)


def test_classifier_matches_reference_golden(cuda):
    p, d = load_golden("kuramoto_clf")
    m = _clf(p, cuda)
    with torch.no_grad():
        feat = m.osc(d["img"].to(cuda))
        logits = m(d["img"].to(cuda))
    assert max_rel(feat, d["feat"]) < RHS_TOL
    assert max_rel(logits, d["logits"]) < RHS_TOL


@pytest.mark.parametrize("H,W,steps", [(28, 28, 10), (5, 7, 3), (1, 1, 2), (32, 32, 1)])
def test_lattice_matches_oracle(cuda, H, W, steps):
    gen = torch.Generator().manual_seed(H * 31 + W)
    x = torch.rand(9, 1, H, W, generator=gen)
    osc = MK.Kuramoto2D(H, W, steps=steps).to(cuda)
    with torch.no_grad():
        osc.omega.copy_(torch.randn(1, 1, H, W, generator=gen) * 0.3)
        osc.K.fill_(0.7)
        got = osc(x.to(cuda))
        want = OL.kuramoto_forward(x, osc.K.cpu(), osc.omega.cpu(), steps=steps, dt=0.15)
    assert got.shape == (9, 2 * H * W)
    assert max_rel(got, want) < RHS_TOL


def test_training_step_gradients_match_oracle_autograd(cuda):
    """fwd+bwd of the whole MNIST model (Kuramoto loop + 1568->10 KAN head + CE loss) against fp64 oracle autograd."""
    p, d = load_golden("kuramoto_clf")
    m = _clf(p, cuda)
    gen = torch.Generator().manual_seed(2)
    img = torch.rand(16, 1, 28, 28, generator=gen)
    labels = torch.randint(0, 10, (16,), generator=gen)
    p64 = {k: v.double().requires_grad_(v.is_floating_point() and "grid" not in k and "neighbor" not in k)
           for k, v in p.items()}
    x64 = img.double().requires_grad_(True)
    feat = OL.kuramoto_forward(x64, p64["osc.K"], p64["osc.omega"])
    head = {k[len("head."):]: v for k, v in p64.items() if k.startswith("head.")}
    loss64 = F.cross_entropy(OL.kanlinear_forward(feat, head), labels)
    loss64.backward()
    xg = img.to(cuda).requires_grad_(True)
    loss = F.cross_entropy(m(xg), labels.to(cuda))
    loss.backward()
    assert abs(float(loss) - float(loss64)) < 1e-5 * max(1.0, abs(float(loss64)))
    assert max_rel(xg.grad, x64.grad) < GRAD_TOL
    for name, t in m.named_parameters():
        assert t.grad is not None, name
        assert max_rel(t.grad, p64[name].grad) < GRAD_TOL, name


def test_empty_batch(cuda):
    osc = MK.Kuramoto2D().to(cuda)
    assert osc(torch.empty(0, 1, 28, 28, device=cuda)).shape == (0, 1568)
