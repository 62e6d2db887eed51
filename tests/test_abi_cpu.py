"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/bt_b200.h declares, and refuses to compute without a GPU (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

import bayesian_torch_b200 as bt
from bayesian_torch_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "bt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bt_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 18
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/bt_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert set(_lib.SIGNATURES) == set(names)


def test_abi_version_and_struct_layout():
    L = _lib.lib()
    assert L.bt_abi_version() == 5
    # struct sizes the header implies (LP64): catches field drift between header and ctypes
    assert ctypes.sizeof(_lib.Geom) == 6 * 4 + 3 * 8 + 6 * 24 + 3 * 8 + 24 + 3 * 8 + 24
    assert ctypes.sizeof(_lib.LayerParams) == 12 * 8 + 2 * 4 + 8 + 2 * 4 + 2 * 8 + 2 * 4 + 8 + 8
    assert ctypes.sizeof(_lib.LayerGrads) == 5 * 8
    assert L.bt_workspace_bytes(b"kl", None, 3) >= 3 * 8 + 8


def test_error_reporting_without_compute():
    L = _lib.lib()
    rc = L.bt_kl_normal_multi(0, None, None, None, 0, 0.0, 1.0, None, None, None, None, None, None, None, None, None)
    assert rc == -1
    assert b"n=0" in L.bt_last_error()
    with pytest.raises(RuntimeError, match="status -1"):
        _lib.check(rc, "kl")


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    layer = bt.layers.LinearReparameterization(6, 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        layer(torch.randn(2, 6))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        layer.kl_loss()
    conv = bt.layers.Conv2dFlipout(3, 4, 3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        conv(torch.randn(1, 3, 8, 8))
