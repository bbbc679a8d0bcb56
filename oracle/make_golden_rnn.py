#!/usr/bin/env python
"""Generate tests/golden/kanrnn_*.npz by EXECUTING THE LIVE REFERENCE KANRNNEncoder / FullyNonlinearKANCell
(train_kan_fet_ett.py:780-818).  TEST INFRASTRUCTURE -- see oracle/__init__.py.  Build container only."""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_live  # noqa: E402
from oracle.make_golden import save  # noqa: E402


def main():
    assert ref_live.available(), "reference checkout missing"
    m = ref_live.script_module("train_kan_fet_ett")
    torch.set_num_threads(1)
    gen = torch.Generator().manual_seed(9)
    # (F, nb, H): the script's sizes (every kept column reads x_t) and a small recurrent case (H > F*nb)
    for tag, F_, nb, H in (("script_7_10_64", 7, 10, 64), ("recurrent_3_4_20", 3, 4, 20)):
        torch.manual_seed(4)
        enc = m.KANRNNEncoder(num_features=F_, hidden_size=H, latent_dim=6, num_basis=nb)
        arrays = {"p." + k: v.detach().clone() for k, v in enc.state_dict().items()}
        x = torch.randn(5, 9, F_, generator=gen)
        h = torch.rand(5, H, generator=gen)
        with torch.no_grad():
            arrays["x"], arrays["h_prev"] = x, h
            arrays["z0"] = enc(x)
            arrays["cell"] = enc.rnn_cell(x[:, 0, :], h)
        save("kanrnn_" + tag, **arrays)


if __name__ == "__main__":
    main()
