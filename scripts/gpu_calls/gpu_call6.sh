#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/f1_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/f1_$name.log | cut -c1-200)"; }
step suite timeout 120 python -m pytest tests -m gpu -q
step bench timeout 60 python bench.py --steps 200 --warmup 20
step convtab timeout 60 python benchmarks/conv_table.py --r50
