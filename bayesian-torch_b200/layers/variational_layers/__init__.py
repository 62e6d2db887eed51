from .linear_variational import *  # noqa: F401,F403
from .conv_variational import *  # noqa: F401,F403
from .rnn_variational import *  # noqa: F401,F403
