#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
N=${NGPU:-2}
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/n${N}_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/n${N}_$name.log | cut -c1-200)"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
step multi timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -s
step trace_c16 timeout 300 $TR --master-port 29611 benchmarks/step_trace.py --comm-ctas 16 --tag c16b
