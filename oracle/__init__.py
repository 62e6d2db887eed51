"""CPU oracle for the Monte-Carlo variational-layer hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` may import it, and only as the checker or the
timed CPU baseline -- never as a fallback for the CUDA path.

Parity status: PINNED against the reference itself.  The reference ships no
tests or golden vectors, so the pins are fixtures made by
importing the reference from ``/root/reference`` and running its own layers on
CPU (``oracle/make_golden.py`` -> ``tests/golden/*.npz``);
``tests/test_oracle_golden.py`` checks every function here against them.
"""
from .bt_oracle import *  # noqa: F401,F403
