#!/usr/bin/env python
"""Secondary measurements (not the driver's bench line): per-layer timings of the wide KANLinear shapes of
BASELINE configs[2..4] on the tcgen05 path vs the CUDA-core path, CUDA-event timed, L2 flushed between iterations.
Prints one JSON object per shape; used to fill profiles/ and DESIGN.md."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "fet-ode_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402

import fetode  # noqa: E402
from fetode._lib import lib  # noqa: E402

SHAPES = [("ecg_kan1", 4096, 768, 128, 10), ("ecg_kan2", 4096, 128, 128, 10), ("ett_enc1", 4096, 2352, 128, 10),
          ("ett_enc1_b64", 64, 2352, 128, 10), ("kanfet_928_256", 4096, 928, 256, 10)]


def timeit(fn, iters=5, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return sorted(ms)[len(ms) // 2]


def main():
    dev = torch.device("cuda:0")
    for name, B, I, O, nb in SHAPES:
        torch.manual_seed(0)
        layer = fetode.KANLinear(I, O, num_basis=nb).to(dev)
        if name.startswith("kanfet"):
            layer.logistic_basis = fetode.TwoDimensionFerroelectricBasis(I, nb).to(dev)
        x = torch.randn(B, I, device=dev)
        out = {}
        with torch.no_grad():
            for mode in ("0", "1"):
                lib().fetode_set_tc_mode(int(mode))
                layer(x)
                out["cuda_core_ms" if mode == "0" else "tcgen05_ms"] = timeit(lambda: layer(x))
        F = 1 + 8 + nb
        flop = 2.0 * B * I * F * O
        out.update(shape=name, B=B, I=I, O=O, nb=nb, useful_gflop=flop / 1e9,
                   tcgen05_useful_tflops=flop / out["tcgen05_ms"] / 1e9, cuda_core_tflops=flop / out["cuda_core_ms"] / 1e9,
                   tcgen05_issued_tf32_tflops=3 * flop / out["tcgen05_ms"] / 1e9)
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
