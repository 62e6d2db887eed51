import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names(prefix=""):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load_golden(name):
    import torch
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = {}
    for k in z.files:
        v = z[k]
        out[k] = torch.from_numpy(v) if v.dtype == np.float32 and not k.startswith("cfg_") else v
    return out


def cfg_of(d):
    """Conv hyper-parameters of a fixture as python ints / tuples."""
    def conv(v):
        v = np.asarray(v)
        return int(v) if v.ndim == 0 else tuple(int(i) for i in v)
    c = {k[4:]: d[k] for k in d if k.startswith("cfg_")}
    out = {k: conv(c[k]) for k in ("stride", "padding", "dilation", "groups", "output_padding") if k in c}
    out["prior_mu"] = float(c.get("prior_mean", 0.0))
    out["prior_sigma"] = float(c.get("prior_variance", 1.0))
    return out


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """The shared library is a build artefact (git-ignored); build it on first use."""
    import importlib.util
    lib = os.path.join(ROOT, "bayesian-torch_b200", "csrc", "libbt_b200.so")
    if not os.path.exists(lib):
        spec = importlib.util.spec_from_file_location("bt_build", os.path.join(ROOT, "bayesian-torch_b200", "csrc", "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()
    return lib
