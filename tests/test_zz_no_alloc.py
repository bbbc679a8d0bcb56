"""fetode_kanlinear_bwd_static: the KANLinear backward with library-owned scratch (a static device array, no allocation on
the call path).  No counterpart in the reference.  Checked through the C
ABI against the existing entry point fetode_kanlinear_bwd (caller-supplied scratch) on two calls in a row that share
the array: a larger layer first, a smaller hysteretic one second, so the second call runs over the first call's
leftovers.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

import fetode
from conftest import max_rel
from fetode._lib import KanLayerGrads, check, lib, ptr, stream

pytestmark = pytest.mark.gpu
TOL = 1e-5               # the bar of the other layer-level parity tests; the two paths run the same kernels, and differ
                         # only in the order of the weight-gradient atomics


def _grads(view, dev):
    """Zeroed gradient tensors for every parameter of the layer + the struct pointing at them."""
    g = KanLayerGrads()
    out = {}
    for name, t in view.tensors():
        if t is None or name == "grid":
            continue
        gt = torch.zeros(t.shape, dtype=torch.float32, device=dev)
        out[name] = gt
        if name.startswith("bp"):
            g.bp[int(name[2:])] = ptr(gt).value
        else:
            setattr(g, name, ptr(gt).value)
    return g, out


def _case(make, B, I, seed, dev, ferro):
    torch.manual_seed(seed)
    layer = make().to(dev)
    s, blob = layer._pack.get()
    x = (torch.rand(B, I, device=dev) * 3 - 1.5).contiguous()
    prev = (torch.rand(B, I, device=dev) * 3 - 1.5).contiguous() if ferro else None
    gy = torch.randn(B, s.out_features, device=dev).contiguous()
    return layer, s, blob, x, prev, gy


def _run_existing(layer, s, blob, x, prev, gy):
    L = lib()
    dev = x.device
    g, outs = _grads(layer._pack.view, dev)
    gx = torch.empty_like(x)
    ws = torch.empty(int(L.fetode_kanlinear_bwd_ws_floats(C.byref(s))), dtype=torch.float32, device=dev)
    check(L.fetode_kanlinear_bwd(C.byref(s), ptr(blob), ptr(x), ptr(prev), C.c_int(1 if prev is None else 0), ptr(gy),
                                 ptr(gx), C.byref(g), ptr(ws), C.c_int64(x.shape[0]), stream()), "kanlinear_bwd")
    torch.cuda.synchronize()
    return gx, outs


def _run_static(layer, s, blob, x, prev, gy):
    L = lib()
    dev = x.device
    g, outs = _grads(layer._pack.view, dev)
    gx = torch.empty_like(x)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0)
    rc = L.fetode_kanlinear_bwd_static(C.byref(s), ptr(blob), ptr(x), ptr(prev), C.c_int(1 if prev is None else 0),
                                       ptr(gy), ptr(gx), C.byref(g), C.c_int64(x.shape[0]), stream())
    torch.cuda.synchronize()
    after = torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0)
    check(rc, "kanlinear_bwd_static")
    assert after == before, "the static-scratch entry point must not allocate through torch"
    return gx, outs


def _compare(got, want, what):
    gx_g, o_g = got
    gx_w, o_w = want
    assert torch.isfinite(gx_g).all()
    assert max_rel(gx_g, gx_w) < TOL, (what, "gx")
    assert set(o_g) == set(o_w)
    for name in ("base_weight", "spline_weight", "branch_weight", "bp0"):
        assert float(o_w[name].abs().max()) > 0, (what, name, "reference gradient is all zero: the case checks nothing")
    for name in o_w:
        assert max_rel(o_g[name], o_w[name]) < TOL, (what, name)


def test_static_scratch_backward_matches_existing_entry_point_twice_in_a_row():
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    L = lib()
    L.fetode_kanlinear_bwd_static_max_floats.restype = C.c_int64
    cap = int(L.fetode_kanlinear_bwd_static_max_floats())
    assert cap == 262144

    # call 1: logistic-branch layer 40 -> 24 (scratch 40*19*24 + 24 floats), 300 rows
    a = _case(lambda: fetode.KANLinear(40, 24), 300, 40, 1, dev, ferro=False)
    need_a = int(L.fetode_kanlinear_bwd_ws_floats(C.byref(a[1])))
    assert 0 < need_a <= cap
    want_a = _run_existing(*a)
    got_a = _run_static(*a)
    _compare(got_a, want_a, "call 1")

    # call 2, right after, on the same array: a smaller hysteretic layer (KANFET's 7 -> 5), 129 rows, other data
    net = fetode.KANFET([7, 5])
    b = _case(lambda: net.layers[0], 129, 7, 2, dev, ferro=True)
    need_b = int(L.fetode_kanlinear_bwd_ws_floats(C.byref(b[1])))
    assert 0 < need_b < need_a
    want_b = _run_existing(*b)
    got_b = _run_static(*b)
    _compare(got_b, want_b, "call 2")

    # and call 1's problem once more: the array now holds call 2's leftovers
    got_a2 = _run_static(*a)
    _compare(got_a2, want_a, "call 3")


def test_static_scratch_backward_rejects_layers_beyond_its_maximum():
    dev = torch.device("cuda:0")
    L = lib()
    layer, s, blob, x, prev, gy = _case(lambda: fetode.KANLinear(256, 128), 4, 256, 3, dev, ferro=False)
    assert int(L.fetode_kanlinear_bwd_ws_floats(C.byref(s))) > 262144
    g, _ = _grads(layer._pack.view, dev)
    gx = torch.full_like(x, 7.0)
    rc = L.fetode_kanlinear_bwd_static(C.byref(s), ptr(blob), ptr(x), None, C.c_int(1), ptr(gy), ptr(gx), C.byref(g),
                                       C.c_int64(4), stream())
    torch.cuda.synchronize()
    assert rc == -5                                                   # FETODE_ETOOLARGE
    assert b"static" in L.fetode_strerror(rc)
    assert bool((gx == 7.0).all()), "a rejected call must not touch its outputs"
