#!/bin/bash
cd "$(dirname "$0")/.."
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']), 'ratio', round(d['e2e']['value']/d['value'],3), 'ms', round(d['ms_per_step'],2))"; }
echo -n "samples=4 :: "; python bench.py --samples 4 --mc-batch 4 --steps 20 --warmup 5 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | pr
echo -n "samples=32 :: "; python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | pr
