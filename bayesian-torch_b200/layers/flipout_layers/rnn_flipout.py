"""LSTMFlipout on the B200-native path: bayesian_torch/layers/flipout_layers/rnn_flipout.py:46-155, the
same unrolled cell as LSTMReparameterization over two LinearFlipout layers."""
from ..variational_layers.rnn_variational import _BayesLSTM
from .linear_flipout import LinearFlipout

__all__ = ["LSTMFlipout"]


class LSTMFlipout(_BayesLSTM):
    _linear = LinearFlipout
