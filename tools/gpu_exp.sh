#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
bash tools/$1.sh > gpurun_out/$1.txt 2>&1; cat gpurun_out/$1.txt
