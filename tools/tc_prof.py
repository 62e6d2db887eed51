"""Cycle accounting of the tensor-core forward for one layer shape (BT_TC2_PROF=1 instantiation).
usage: tc_prof.py cin cout k stride pad batch hw [samples]   (same arguments as one_layer.py)
Prints, per warp role, the mean over CTAs of the cycles spent in each kind of wait."""
import ctypes as C
import os
import subprocess
import sys

os.environ["BT_TC2_PROF"] = "1"
os.environ.setdefault("BT_AFF", "1")
os.environ.setdefault("BT_XS", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.argv = [sys.argv[0], *sys.argv[1:9], "1"] if len(sys.argv) > 8 else [*sys.argv, "1", "1"]
exec(open(os.path.join(ROOT, "tools", "one_layer.py")).read())  # runs the layer (iters = 1)
import numpy as np  # noqa: E402
from bayesian_torch_b200 import _lib  # noqa: E402

buf = np.zeros(160 * 6 * 8, dtype=np.uint64)
rc = _lib.lib().bt_debug_tc_profile(buf.ctypes.data_as(C.c_void_p), buf.size)
assert rc == 0
p = buf.reshape(160, 6, 8).astype(np.float64)
live = p[:, 0, 7] > 0
p = p[live]
names = {
    0: ("epilogue", ["wait acc_full", "wait residual", "tmem ld wait", "store drain", "convert+stage", "fence+store issue", "residual issue", "total"]),
    1: ("tma loader", ["wait a_empty", "-", "-", "-", "-", "-", "-", "total"]),
    2: ("mma issuer 0", ["wait acc_empty", "wait w_full", "wait a_full", "-", "-", "-", "-", "total"]),
    3: ("mma issuer 1", ["wait acc_empty", "wait w_full", "wait a_full", "-", "-", "-", "-", "total"]),
    4: ("producer warp 0", ["wait w_empty", "generate", "-", "-", "-", "-", "-", "total"]),
    5: ("producer warp 15", ["wait w_empty", "generate", "-", "-", "-", "-", "-", "total"]),
}
print(f"{int(live.sum())} CTAs; mean cycles per CTA (max over CTAs in brackets)")
for r, (role, slots) in names.items():
    tot = p[:, r, 7].mean()
    parts = ", ".join(f"{n} {p[:, r, i].mean():.0f} ({100 * p[:, r, i].mean() / max(tot, 1):.0f}%)"
                      for i, n in enumerate(slots[:-1]) if n != "-")
    print(f"  {role:17s} total {tot:.0f} [{p[:, r, 7].max():.0f}]  {parts}")
