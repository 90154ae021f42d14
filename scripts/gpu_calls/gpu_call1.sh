#!/bin/bash
# Round-2 first GPU call (1 GPU): default suite + bench, then each never-executed variant in isolation.
# Every step is bounded by its own timeout and logs to gpurun_out/c1_*.log.
set -u
O=gpurun_out
mkdir -p $O
NAT=distributed-pytorch-training_b200/_native
step() { local name=$1; shift; local t0=$(date +%s); "$@" > $O/c1_$name.log 2>&1; local rc=$?
         echo "[$name] rc=$rc $(( $(date +%s) - t0 ))s: $(tail -1 $O/c1_$name.log | cut -c1-300)"; }

nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $O/c1_clocks.csv &
SMI=$!

step suite_default timeout 900 python -m pytest tests -m gpu -x -q
step bench_default timeout 300 python bench.py --steps 200 --warmup 20
step r50_default timeout 400 python bench.py --arch resnet50 --batch-size 256 --image-size 224 --num-classes 1000 --steps 10 --warmup 3

# 2: 128-bit vectors
cp $NAT/libdptb200.so /tmp/libdptb200.default.so
cp build/variants/libdptb200.vec128.so $NAT/libdptb200.so
step suite_vec128 timeout 900 python -m pytest tests -m gpu -q
step bench_vec128 timeout 300 python bench.py --steps 200 --warmup 20
step r50_vec128 timeout 400 python bench.py --arch resnet50 --batch-size 256 --image-size 224 --num-classes 1000 --steps 10 --warmup 3
# 3: bulk BN kernels + large shapes, three processes (still on the vec128 build)
for i in 1 2 3; do
  DPT_BULK_EW=1 DPT_TEST_LARGE=1 step bulk_large_$i timeout 300 python -m pytest tests/test_gpu_ops.py tests/test_gpu_engine.py -k "large or imagenet" -x -q
done
DPT_BULK_EW=1 step r50_vec128_bulk timeout 400 python bench.py --arch resnet50 --batch-size 256 --image-size 224 --num-classes 1000 --steps 10 --warmup 3
cp /tmp/libdptb200.default.so $NAT/libdptb200.so
for i in 1 2; do
  DPT_BULK_EW=1 DPT_TEST_LARGE=1 step bulk_large_default_$i timeout 300 python -m pytest tests/test_gpu_ops.py tests/test_gpu_engine.py -k "large or imagenet" -x -q
done

step kernel_bench timeout 300 python benchmarks/kernel_bench.py
# 4/5: 128x256 tiles, two-SM GEMM (a wrong barrier count hangs: short timeouts)
DPT_TEST_BN256=1 step bn256 timeout 90 python -m pytest tests/test_gpu_igemm.py -k block_n_256 -x -q
DPT_TEST_2SM=1 step gemm2sm timeout 90 python -m pytest tests/test_gpu_gemm2sm.py -x -q
kill $SMI
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm --format=csv > $O/c1_final_smi.csv 2>&1
echo done
