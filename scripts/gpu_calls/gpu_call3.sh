#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/c6_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/c6_$name.log | cut -c1-400)"; }
step suite timeout 900 python -m pytest tests -m gpu -x -q
step bench timeout 300 python bench.py --steps 200 --warmup 20
step r50 timeout 400 python bench.py --arch resnet50 --batch-size 256 --image-size 224 --num-classes 1000 --steps 10 --warmup 3
step trace1 timeout 300 python benchmarks/step_trace.py
for i in 1 2; do step large_$i timeout 300 python -m pytest tests/test_gpu_ops.py tests/test_gpu_engine.py -k "large or imagenet" -x -q; done
