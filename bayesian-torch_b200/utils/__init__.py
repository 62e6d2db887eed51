from .util import get_rho  # noqa: F401
