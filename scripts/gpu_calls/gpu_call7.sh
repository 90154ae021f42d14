#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 40 python -m pytest tests/test_gpu_multi.py -q -s -k "checkpoint" > $O/f4_multi.log 2>&1
echo "rc=$? $(tail -1 $O/f4_multi.log | cut -c1-200)"
