"""bayesian_torch_b200 -- B200-native Monte-Carlo variational-layer hot path.

Drop-in for the layer / dnn_to_bnn surface of godhj93/bayesian-torch:
    from bayesian_torch_b200.layers import LinearReparameterization, Conv2dFlipout, ...
    from bayesian_torch_b200.models.dnn_to_bnn import dnn_to_bnn, get_kl_loss
Everything computes in hand-written sm_100a kernels behind the C ABI of
include/bt_b200.h (libbt_b200.so); there is no CPU or eager fallback.

Layout: this directory is the importable package (a hyphen is not a valid identifier); its
sub-modules and the CUDA sources live next door in `bayesian-torch_b200/`, the directory name
the build layout prescribes, which is appended to the package search path below -- the
standard `__path__` mechanism (what pkgutil.extend_path does), so `inspect`, pickling and
`multiprocessing` spawn see ordinary modules named bayesian_torch_b200.*.
"""
import os as _os

__path__.append(_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "bayesian-torch_b200"))

from . import _lib, layers, models, ops, utils  # noqa: E402,F401
from .mc import (GradientBuckets, HostBatchPrefetcher, MCAccumulator, MCPredictor, MCResult, allreduce_gradients,  # noqa: E402,F401
                 allreduce_moments, mc_predict, shard_samples)
from .models.dnn_to_bnn import dnn_to_bnn, get_kl_loss  # noqa: E402,F401
from .models.fuse import fold_batchnorm  # noqa: E402,F401
from .ops import (invalidate_packed, manual_seed, mc_batched, mc_sample, path, sample_mode, set_path,  # noqa: E402,F401
                  set_sample_mode)

__version__ = "0.2.0"
