#!/bin/bash
# multi-GPU call: multi-GPU tests, step traces (comm CTA sweep), bench
set -u
O=gpurun_out; mkdir -p $O
N=${NGPU:-2}
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/m${N}_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/m${N}_$name.log | cut -c1-260)"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
step multi timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -s
P=29601
for c in 16 32 48 64; do
  step trace_c$c timeout 300 $TR --master-port $P benchmarks/step_trace.py --comm-ctas $c --tag c$c; P=$((P+1))
done
step trace_strict timeout 300 $TR --master-port $P benchmarks/step_trace.py --comm-ctas 16 --strict-barriers --tag strict16; P=$((P+1))
step bench timeout 300 $TR --master-port $P bench.py --gpus $N --steps 200 --warmup 20
step bench1 timeout 300 python bench.py --steps 200 --warmup 20 --no-extras
