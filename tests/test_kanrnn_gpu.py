"""GPU parity of the one-launch logistic-KAN recurrent encoder (csrc/kanrnn.cu) through the C ABI: golden vectors
from the live reference (oracle/make_golden_rnn.py), gradients against fp64 oracle autograd in both regimes (the
script's sizes, where h_T depends on the last context row only, and H > F*nb, where the recurrence is live), and the
script's full context length."""
import pytest
import torch

import fetode
from conftest import load_golden, max_rel
from fetode import ett
from oracle import layers as OL

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag,F,nb,H", [("script_7_10_64", 7, 10, 64), ("recurrent_3_4_20", 3, 4, 20)])
def test_encoder_and_cell_golden(cuda, tag, F, nb, H):
    p, d = load_golden("kanrnn_" + tag)
    enc = ett.KANRNNEncoder(num_features=F, hidden_size=H, latent_dim=6, num_basis=nb)
    enc.load_state_dict(p)
    enc = enc.to(cuda)
    with torch.no_grad():
        assert max_rel(enc(d["x"].to(cuda)), d["z0"]) < 1e-5
        assert max_rel(enc.rnn_cell(d["x"][:, 0, :].to(cuda), d["h_prev"].to(cuda)), d["cell"]) < 1e-5


@pytest.mark.parametrize("F,nb,H,T,B", [(7, 10, 64, 336, 64), (3, 4, 20, 17, 33), (2, 2, 40, 5, 7)])
def test_encoder_gradients_vs_fp64_oracle(cuda, F, nb, H, T, B):
    torch.manual_seed(F * 100 + H)
    enc = ett.KANRNNEncoder(num_features=F, hidden_size=H, latent_dim=8, num_basis=nb)
    p = {k: v.detach().clone() for k, v in enc.state_dict().items()}
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    enc = enc.to(cuda)
    gen = torch.Generator().manual_seed(T)
    x = torch.randn(B, T, F, generator=gen)
    w = torch.randn(B, 8, generator=gen, dtype=torch.float64)
    x64 = x.double().requires_grad_(True)
    z64 = OL.kan_rnn_encoder(x64, p64)
    (z64 * w).sum().backward()
    xg = x.to(cuda).requires_grad_(True)
    z = enc(xg)
    assert max_rel(z, z64) < 1e-5
    (z * w.float().to(cuda)).sum().backward()
    assert max_rel(xg.grad, x64.grad) < 2e-4
    for name, t in enc.named_parameters():
        g64 = p64[name].grad
        if g64 is None or float(g64.abs().max()) == 0.0:           # hidden_basis is dead when every column reads x_t
            assert t.grad is None or float(t.grad.abs().max()) == 0.0, name
            continue
        assert max_rel(t.grad, g64) < 2e-4, name


def test_cell_with_initial_state_gradient(cuda):
    torch.manual_seed(3)
    cell = ett.FullyNonlinearKANCell(2, 30, 3)
    p = {k: v.detach().clone() for k, v in cell.state_dict().items()}
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    cell = cell.to(cuda)
    gen = torch.Generator().manual_seed(1)
    x, h = torch.randn(9, 2, generator=gen), torch.rand(9, 30, generator=gen)
    x64, h64 = x.double().requires_grad_(True), h.double().requires_grad_(True)
    OL.kan_rnn_cell(x64, h64, p64).square().sum().backward()
    xg, hg = x.to(cuda).requires_grad_(True), h.to(cuda).requires_grad_(True)
    cell(xg, hg).square().sum().backward()
    assert max_rel(xg.grad, x64.grad) < 2e-4 and max_rel(hg.grad, h64.grad) < 2e-4
