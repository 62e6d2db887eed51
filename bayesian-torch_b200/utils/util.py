"""get_rho / MC uncertainty helpers on the GPU path (reference: bayesian_torch/utils/util.py)."""
import torch

from .. import ops


def get_rho(sigma, delta):
    """MOPED: rho = log(expm1(delta*|w|) + 1e-20) (utils/util.py:63-69).  CUDA tensors run the
    bt_moped_rho kernel; dnn_to_bnn() calls this on the CPU-resident DNN weights before the
    model is moved to the GPU, so CPU tensors are initialised with the same formula in torch
    (one-off host-side init, not part of the hot path)."""
    if sigma.is_cuda and sigma.dtype == torch.float32:
        return ops.moped_rho(sigma, delta)
    return torch.log(torch.expm1(delta * torch.abs(sigma)) + 1e-20)
