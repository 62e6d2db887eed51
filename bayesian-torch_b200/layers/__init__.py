"""Flat re-export, as bayesian_torch/layers/__init__.py:1-6 does: dnn_to_bnn resolves
classes by name with getattr(bayesian_layers, "<TorchClass><Type>")."""
from .base_variational_layer import BaseVariationalLayer_, get_kernel_size  # noqa: F401
from .variational_layers import *  # noqa: F401,F403
from .flipout_layers import *  # noqa: F401,F403
