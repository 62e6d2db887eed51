"""Per-layer timing table of the Bayesian ResNet-50 forward (one weight sample, batch 128):
CUDA events around every layer call -> ms, TFLOP/s, GB/s per distinct layer shape."""
import collections
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import ops  # noqa: E402
from bench import build_model  # noqa: E402

dev = torch.device("cuda:0")
mc_batch = int(sys.argv[1]) if len(sys.argv) > 1 else 1
model = build_model(dev, fold_bn=os.environ.get('BT_FOLD', '1') == '1')
x = torch.randn(128, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
names = {}
for n, m in model.named_modules():
    if hasattr(m, "_layer_id"):
        names[m._layer_id] = (n, m)


def run():
    with torch.no_grad():
        if mc_batch == 1:
            with bt.mc_sample(0):
                model(x)
        else:
            with bt.mc_batched(mc_batch, 0, shared_batch=128):
                model(x)


for _ in range(2):
    run()
torch.cuda.synchronize()
ops.PROFILE.reset(); ops.PROFILE.timing = True
order = []
orig = ops._BayesLayerFn.forward
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record()
torch.cuda.synchronize()
ops.PROFILE.timing = False
total = e0.elapsed_time(e1)
layers = [m for _, m in sorted(names.items())]
rows = collections.OrderedDict()
for (path, flops, nbytes, a, b), (lid, (name, m)) in zip(ops.PROFILE.records, sorted(names.items())):
    w = m._w()[0]
    key = (tuple(w.shape), m.stride if hasattr(m, "stride") else 1, flops, nbytes)   # (nbytes: with / without a residual input)
    r = rows.setdefault(key, {"name": name, "n": 0, "ms": 0.0, "flops": flops, "bytes": nbytes, "path": path})
    r["n"] += 1
    r["ms"] += a.elapsed_time(b)
print(f"mc_batch={mc_batch} whole forward {total:.2f} ms; sum of Bayesian layer calls {sum(r['ms'] for r in rows.values()):.2f} ms")
PEAK_TF, PEAK_GB = float(os.environ.get("BT_PEAK_TF", "1372.6")), float(os.environ.get("BT_PEAK_GB", "6541.5"))
print(f"{'first layer':28s} {'weight':22s} {'cnt':>3s} {'ms/call':>8s} {'TFLOP/s':>8s} {'GB/s':>7s} {'t_tc':>6s} {'t_hbm':>6s} {'gap ms':>7s} path")
gap_total = ideal_total = 0.0
for key, r in rows.items():
    ms = r["ms"] / r["n"]
    t_tc, t_hbm = r["flops"] / PEAK_TF / 1e9, r["bytes"] / PEAK_GB / 1e6
    gap = (ms - max(t_tc, t_hbm)) * r["n"]
    gap_total += gap; ideal_total += max(t_tc, t_hbm) * r["n"]
    print(f"{r['name']:28s} {str(key[0]):22s} {r['n']:3d} {ms:8.3f} {r['flops'] / ms / 1e9:8.1f} {r['bytes'] / ms / 1e6:7.0f} "
          f"{t_tc:6.3f} {t_hbm:6.3f} {gap:7.3f} {'tc' if r['path'] == 2 else 'simt'}")
print(f"sum of per-layer max(t_tc, t_hbm) = {ideal_total:.2f} ms; gap to it {gap_total:.2f} ms (peaks {PEAK_TF} TFLOP/s, {PEAK_GB} GB/s)")
