#!/bin/bash
# 8-GPU call: multi-GPU tests at world 8, step traces, scaling benches
set -u
O=gpurun_out; mkdir -p $O
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/g8_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/g8_$name.log | cut -c1-200)"; }
TRUN="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
step multi timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -s
step trace_c16 timeout 300 $TRUN --nproc-per-node 8 --master-port 29621 benchmarks/step_trace.py --comm-ctas 16 --tag c16
step trace_c32 timeout 300 $TRUN --nproc-per-node 8 --master-port 29622 benchmarks/step_trace.py --comm-ctas 32 --tag c32
step trace_mc timeout 300 $TRUN --nproc-per-node 8 --master-port 29623 benchmarks/step_trace.py --comm-ctas 16 --allreduce-algo twoshot_mc --tag mc16
step bench8 timeout 300 $TRUN --nproc-per-node 8 --master-port 29624 bench.py --gpus 8 --steps 200 --warmup 20
step bench8_c32 timeout 300 $TRUN --nproc-per-node 8 --master-port 29625 bench.py --gpus 8 --steps 200 --warmup 20 --comm-ctas 32 --no-extras
step bench4 timeout 300 $TRUN --nproc-per-node 4 --master-port 29626 bench.py --gpus 4 --steps 200 --warmup 20 --no-extras
step bench2 timeout 300 $TRUN --nproc-per-node 2 --master-port 29627 bench.py --gpus 2 --steps 200 --warmup 20 --no-extras
step bench1 timeout 300 python bench.py --steps 200 --warmup 20 --no-extras
