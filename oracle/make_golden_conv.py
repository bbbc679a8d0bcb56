#!/usr/bin/env python
"""Generate tests/golden/ferroconv_*.npz by EXECUTING THE LIVE REFERENCE classes FerroelectricBasisConv2d and
MemEfficient_FerroelectricBasisConv2d (ferro_class.py:601-944) on seeded inputs.  TEST INFRASTRUCTURE -- see
oracle/__init__.py.  Build container only:   python oracle/make_golden_conv.py"""
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
    fc = ref_live.ferro_class()
    torch.set_num_threads(1)
    gen = torch.Generator().manual_seed(77)
    for tag, cls, cout, kw in (
            ("plain_s1p1", fc.FerroelectricBasisConv2d, 5, dict(kernel_size=3, stride=1, padding=1)),
            ("plain_s2p0_c1", fc.FerroelectricBasisConv2d, 1, dict(kernel_size=(3, 2), stride=2, padding=0)),
            ("memeff_s1p1", fc.MemEfficient_FerroelectricBasisConv2d, 5, dict(kernel_size=3, stride=1, padding=1, out_chunk=2))):
        torch.manual_seed(5)
        m = cls(2, cout, num_basis=3, **kw)
        with torch.no_grad():
            m.out_bias.copy_(torch.randn(cout, generator=gen) * 0.3)
        arrays = {"p." + k: v.detach().clone() for k, v in m.state_dict().items() if not k.startswith("_")}
        xs = [torch.randn(3, 2, 7, 6, generator=gen) * 1.5 for _ in range(3)]
        with torch.no_grad():
            arrays["x0"], arrays["x1"], arrays["x2"] = xs
            arrays["stateless0"] = m(xs[0])                       # default mode: every call independent
            arrays["stateless1"] = m(xs[1])
            m.stateful = True
            m.reset_state()
            if cout == 1 or "MemEfficient" in cls.__name__:       # the plain class's stateful mode needs Cout == 1
                for i, x in enumerate(xs):                        # true hysteresis across calls
                    arrays[f"stateful{i}"] = m(x)
            else:
                try:
                    m(xs[0])
                    raise AssertionError("expected the reference to fail")
                except RuntimeError:
                    arrays["stateful_raises"] = torch.ones(1)
        save("ferroconv_" + tag, **arrays)


if __name__ == "__main__":
    main()
