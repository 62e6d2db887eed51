"""bayesian_torch_b200 -- B200-native Monte-Carlo variational-layer hot path.

Drop-in for the layer / dnn_to_bnn surface of godhj93/bayesian-torch:
    from bayesian_torch_b200.layers import LinearReparameterization, Conv2dFlipout, ...
    from bayesian_torch_b200.models.dnn_to_bnn import dnn_to_bnn, get_kl_loss
Everything computes in hand-written sm_100a kernels behind the C ABI of
include/bt_b200.h (libbt_b200.so); there is no CPU or eager fallback.
"""
from . import _lib, layers, models, ops, utils  # noqa: F401
from .mc import (GradientBuckets, HostBatchPrefetcher, MCAccumulator, MCPredictor, MCResult, allreduce_gradients,  # noqa: F401
                 allreduce_moments, mc_predict, shard_samples)  # noqa: F401
from .models.dnn_to_bnn import dnn_to_bnn, get_kl_loss  # noqa: F401
from .models.fuse import fold_batchnorm  # noqa: F401
from .ops import (invalidate_packed, manual_seed, mc_batched, mc_sample, path, sample_mode, set_path,  # noqa: F401
                  set_sample_mode)

__version__ = "0.1.0"
