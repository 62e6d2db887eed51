"""Device timeline of one bench step (torch.profiler / CUPTI): busy time, idle gaps between consecutive kernels,
and the largest gaps with the kernels on either side.  usage: gap_profile.py [samples]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
import bench  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda:0")
model = bench.build_model(dev)
x = torch.randn(128, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
for _ in range(3):
    bt.mc_predict(model, x, S, batch_samples=S)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        bt.mc_predict(model, x, S, batch_samples=S)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
ev.sort(key=lambda e: e.time_range.start)
ev = ev[len(ev) // 2:]          # second step only
busy = sum(e.time_range.end - e.time_range.start for e in ev)
span = ev[-1].time_range.end - ev[0].time_range.start
gaps = [(b.time_range.start - a.time_range.end, a.name[:60], b.name[:60]) for a, b in zip(ev, ev[1:])]
print(f"{len(ev)} kernels, span {span / 1e3:.2f} ms, busy {busy / 1e3:.2f} ms, idle {sum(g for g, _, _ in gaps) / 1e3:.2f} ms")
by = {}
for g, a, b in gaps:
    k = (a.split("<")[0][-40:], b.split("<")[0][-40:])
    by.setdefault(k, [0, 0.0]); by[k][0] += 1; by[k][1] += g
for k, (n, t) in sorted(by.items(), key=lambda kv: -kv[1][1])[:12]:
    print(f"  {t / 1e3:7.3f} ms over {n:3d} gaps (avg {t / n:6.1f} us)  {k[0]} -> {k[1]}")
tot = {}
for e in ev:
    k = e.name.split("(")[0][-60:]
    tot.setdefault(k, [0, 0.0]); tot[k][0] += 1; tot[k][1] += e.time_range.end - e.time_range.start
for k, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:8]:
    print(f"  {t / 1e3:7.3f} ms {n:3d}x {k}")
