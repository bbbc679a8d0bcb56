"""Host-side state handling of the stateful neurons (no kernel is launched): the registered ``prev_x`` buffer is a
stride-0 view of compact rows after any forward, and a checkpoint restore (train_kan_fet_ett.py:353-358 reloads the
best ``state_dict``, buffers included) must still work -- the reference keeps a real [B,I,(O,)K] buffer there."""
import copy

import torch

import fetode
from fetode import ecg, ferro_class as FC


def test_two_d_state_dict_round_trip_after_forward_state():
    m = FC.TwoDimensionFerroelectricBasis(3, 4)
    rows = torch.randn(5, 3)
    m._set_rows(rows.clone())
    assert m.prev_x.shape == (5, 3, 4)
    sd = copy.deepcopy(m.state_dict())
    m.load_state_dict(sd)                                       # used to raise: copy_ into an overlapping view
    assert torch.equal(m._rows, rows) and torch.equal(m.prev_x, rows[:, :, None].expand(5, 3, 4))
    fresh = FC.TwoDimensionFerroelectricBasis(3, 4)
    fresh.load_state_dict(sd)
    assert torch.equal(fresh._rows, rows)
    fresh.prev_x.zero_()                                        # the reference's in-place reset still reaches the state
    assert not fresh._rows.any()


def test_full_ferro_state_dict_round_trip_and_fresh_buffers():
    m = FC.NoisyFerroelectricBasis(3, 2, 4)
    sd0 = copy.deepcopy(m.state_dict())
    m.load_state_dict(sd0)
    assert m._rows is None and m.prev_x.shape == (1, 3, 2, 4)   # fresh [1,...] zeros stay "fresh"
    rows = torch.randn(6, 3)
    m._set_rows(rows.clone())
    m.branch_sign = torch.rand(6, 3, 2, 4)
    m._br_live = True
    sd = copy.deepcopy(m.state_dict())
    other = FC.NoisyFerroelectricBasis(3, 2, 4)
    other.load_state_dict(sd)
    assert torch.equal(other._rows, rows) and other._br_live
    assert torch.equal(other.branch_sign, sd["branch_sign"])


def test_kanfet_checkpoint_restore_with_live_state():
    torch.manual_seed(0)
    net = fetode.KANFET([2, 10, 2])
    for l in net.layers:
        l.logistic_basis._set_rows(torch.randn(7, l.in_features))
    best = copy.deepcopy(net.state_dict())
    for l in net.layers:
        l.logistic_basis._set_rows(torch.randn(7, l.in_features))
    net.load_state_dict(best)
    for i, l in enumerate(net.layers):
        assert torch.equal(l.logistic_basis.prev_x, best[f"layers.{i}.logistic_basis.prev_x"])


def test_hardgate_basis_state_dict_round_trip():
    m = ecg.LogisticBasis(4, 3)
    row = torch.randn(4)
    m._row = row.clone()
    m.prev_x = m._row.view(1, -1, 1).expand(1, 4, 3)
    m.branch_state = torch.ones(5, 4, 3, dtype=torch.uint8)
    sd = copy.deepcopy(m.state_dict())
    m.load_state_dict(sd)
    assert torch.equal(m._prev_row(torch.device("cpu")), row)
    other = ecg.LogisticBasis(4, 3)
    other.load_state_dict(sd)
    assert torch.equal(other._prev_row(torch.device("cpu")), row)
