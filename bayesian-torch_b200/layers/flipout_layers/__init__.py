from .linear_flipout import *  # noqa: F401,F403
from .conv_flipout import *  # noqa: F401,F403
from .rnn_flipout import *  # noqa: F401,F403
