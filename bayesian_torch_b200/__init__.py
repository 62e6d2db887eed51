"""Importable alias of the package directory `bayesian-torch_b200/` (a hyphen is not a
valid Python identifier).  The real sources live there; this module only redirects."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "bayesian-torch_b200"))
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _os, _f
