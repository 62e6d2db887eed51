#!/bin/bash
cd "$(dirname "$0")/.."
python tools/e2e_probe.py e2e 2>&1 | tail -6
python tools/e2e_probe.py resident 2>&1 | tail -6
BT_PDL=0 python tools/e2e_probe.py e2e 2>&1 | tail -6
