#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
{
DPT_BULK_EW=4 ITERS=6 timeout 120 python scripts/diag_bulk.py 2>&1 | grep -v "^    vec\|vs a"
DPT_BULK_EW=7 ITERS=6 RELU=1 timeout 120 python scripts/diag_bulk.py 2>&1 | grep -v "^    vec\|vs a"
} > $O/c5.log 2>&1
cat $O/c5.log
