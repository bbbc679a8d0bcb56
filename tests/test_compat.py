"""fetode.compat: the reference's own import lines resolve to the drop-in ('scripts run unchanged
after an import swap').  The second half needs the reference checkout (build container): it imports two reference
SCRIPTS through the shim, builds their models, and checks that parameter / buffer names and shapes equal those of the
same scripts imported against the reference's own modules -- i.e. ``state_dict`` files are interchangeable."""
import importlib.util
import os
import sys

import pytest
import torch

import fetode
from fetode import compat
from oracle import ref_live

_NAMES = ("efficientkan", "efficient_kan", "efficient_kan.efficientkan", "torchdiffeq", "ferro_class",
          "kan_diffusion", "kan_diffusion.kan")


@pytest.fixture
def shim():
    saved = {n: sys.modules.get(n) for n in _NAMES}
    yield lambda **kw: compat.install(force_torchdiffeq=True, **kw)
    compat.uninstall()
    for n, m in saved.items():
        if m is not None:
            sys.modules[n] = m
        else:
            sys.modules.pop(n, None)


def test_reference_import_lines_resolve_to_fetode(shim):
    names = shim()
    assert set(names) == set(_NAMES)
    import efficientkan                                          # train_kanfet_node_predprey.py:12
    from efficient_kan.efficientkan import KANFET                # train_ecg.py:16
    from ferro_class import BatchedFerroelectricBasis, FerroelectricBasis     # train_ecg.py:17
    from kan_diffusion import kan                                # train_ecg_kan_fet_nn_ode.py:14
    from kan_diffusion.kan import KAN                            # train_kan_fet_ett.py:13
    from torchdiffeq import odeint                               # train_kanfet_node_predprey.py:5
    assert efficientkan.KANFET is fetode.KANFET is KANFET
    assert FerroelectricBasis is fetode.FerroelectricBasis and BatchedFerroelectricBasis.__module__ == "fetode.ferro_class"
    assert kan.KAN is KAN is fetode.KAN
    assert odeint is fetode.odeint
    net = efficientkan.KANFET(layers_hidden=[2, 10, 2], grid_size=5)          # the script's constructor call :146
    assert sum(p.numel() for p in net.parameters()) == 1412


def _load_script(name, alias):
    path = os.path.join(ref_live.REF_ROOT, name + ".py")
    spec = importlib.util.spec_from_file_location(alias, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _shapes(m):
    return {k: tuple(v.shape) for k, v in m.state_dict().items()}


@pytest.mark.skipif(not ref_live.available(), reason="reference checkout not present")
def test_reference_scripts_build_the_same_state_dict_through_the_shim(shim):
    ref_a = ref_live.script_module("train_ecg_kan_fet_nn_ode")    # reference modules (+ stubs for absent packages)
    ref_b = ref_live.script_module("train_ecg")
    torch.manual_seed(0)
    want_a = _shapes(ref_a.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128))
    want_b = _shapes(ref_b.KANFetODEFunc(latent_dim=16, hidden_dim=24, num_basis=6))
    shim()
    if ref_live.REF_ROOT not in sys.path:
        sys.path.insert(0, ref_live.REF_ROOT)
    a = _load_script("train_ecg_kan_fet_nn_ode", "shim_train_ecg_kan_fet_nn_ode")
    b = _load_script("train_ecg", "shim_train_ecg")
    assert a.odeint is fetode.odeint and a.kan.KAN is fetode.KAN
    assert b.FerroelectricBasis is fetode.FerroelectricBasis and b.KANFET is fetode.KANFET
    model_a = a.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128)
    assert isinstance(model_a.odefunc.kan1, fetode.KAN)
    assert _shapes(model_a) == want_a
    model_b = b.KANFetODEFunc(latent_dim=16, hidden_dim=24, num_basis=6)
    assert isinstance(model_b.fc1, fetode.FerroelectricBasis)
    assert _shapes(model_b) == want_b
    # and the reference's own checkpoints load into the shimmed model
    torch.manual_seed(1)
    src = ref_a.KanFet_MLP_NODE(T=96, num_classes=2, latent_dim=64, num_basis=12, ode_hidden=128)
    model_a.load_state_dict(src.state_dict())
