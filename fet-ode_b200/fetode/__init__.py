"""fetode -- B200 (sm_100a) implementation of the KAN-FET neural-ODE right-hand side.

Host-side mirror of the reference's module surface (sallywang147/FET-ODE @ 704a0df4):
  fetode.efficientkan   LogisticBasis, KANLinear, KAN, KANFET      (efficient_kan/efficientkan.py)
  fetode.ferro_class    *FerroelectricBasis                         (ferro_class.py)
  fetode.ecg            hysteretic LogisticBasis, KANFeatureMixer, MLPKANODEFunc, KANFetODEFunc (ECG scripts)
  fetode.mnist_kuramoto_kan  KANLinear, Kuramoto2D, KuramotoKANClassifier             (mnist_kuramoto_kan.py)
  fetode.ett            LogisticBasisLinear, FullyNonlinearKANCell, KANRNNEncoder     (train_kan_fet_ett.py:741-818)
  fetode.diffusion      p_sample_loop over the KANFET denoiser, graph-replayed        (kan_diffusion_ett.py:565-609)
  fetode.odeint         odeint, integrate_euler/rk2/rk4, odeint_rk4 (torchdiffeq call sites + in-repo loops)
  fetode.compat         install(): sys.modules shims so the reference's own import lines resolve to the above
All arithmetic runs in fetode/libfetode.so (C ABI: include/fetode.h); there is no CPU fallback.
"""
from . import _lib  # noqa: F401
from .efficientkan import KAN, KANFET, KANLinear, LogisticBasis  # noqa: F401
from .ferro_class import (BatchedFerroelectricBasis, FerroelectricBasis, FerroelectricBasisConv2d,  # noqa: F401
                          MemEfficient_FerroelectricBasisConv2d, NoisyBatchedFerroelectricBasis,
                          OriginalBatchedFerroelectricBasis,
                          NoisyFerroelectricBasis, TwoDimensionFerroelectricBasis)
from . import compat, diffusion, ecg, ett, graphs, mnist_kuramoto_kan, optim  # noqa: F401
from .odeint import (as_rhs, integrate_euler, integrate_rk2, integrate_rk4, odeint,  # noqa: F401
                     odeint_rk4)

__all__ = ["KAN", "KANFET", "KANLinear", "LogisticBasis", "TwoDimensionFerroelectricBasis", "FerroelectricBasis",
           "NoisyFerroelectricBasis", "BatchedFerroelectricBasis", "NoisyBatchedFerroelectricBasis",
           "FerroelectricBasisConv2d", "MemEfficient_FerroelectricBasisConv2d", "OriginalBatchedFerroelectricBasis", "ecg",
           "mnist_kuramoto_kan", "optim", "graphs", "compat", "ett", "diffusion", "odeint",
           "integrate_euler", "integrate_rk2", "integrate_rk4", "odeint_rk4", "as_rhs"]
