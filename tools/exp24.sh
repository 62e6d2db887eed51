#!/bin/bash
# launch list of one ResNet-18 Flipout training step (bench.py --workload train)
cd "$(dirname "$0")/.."
python bench.py --workload train --steps 1 --warmup 1 > gpurun_out/train_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_train_launches_raw.csv \
  python bench.py --workload train --steps 1 --warmup 1 > gpurun_out/ncu_train.log 2>&1
echo "rc=$?"
python - <<'PY'
import csv, collections
lines = [l for l in open('gpurun_out/r2_train_launches_raw.csv') if not l.startswith('==')]
rows = list(csv.DictReader(lines))
n = len(rows)
# the last third of the launches ~ the last (e2e) steps; summarise the final 1/6 as one step
agg = collections.defaultdict(float); cnt = collections.Counter()
tail = rows[-(n // 8):]
for r in tail:
    name = r['Kernel Name'].split('(')[0].replace('void ', '')[-60:]
    v = float(r['Metric Value'].replace(',', '')); u = r['Metric Unit']
    v *= {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0}.get(u, 1e-6)
    agg[name] += v; cnt[name] += 1
tot = sum(agg.values())
print(n, 'launches in all;', len(tail), 'in the tail window,', round(tot, 2), 'ms')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:22]:
    print(f'{v:8.3f} ms {100 * v / tot:5.1f}% {cnt[k]:5d}  {k}')
PY
