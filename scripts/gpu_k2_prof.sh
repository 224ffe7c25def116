#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/bench_attention.py --impls 0 --only single --iters 5"
$CMD > gpurun_out/k2_plain.log 2>&1 || { tail gpurun_out/k2_plain.log; exit 1; }
tail -1 gpurun_out/k2_plain.log
ncu --set full --clock-control none --import-source on -k regex:attn_v1 -s 4 -c 1 -o gpurun_out/k2_v2 -f $CMD > gpurun_out/ncu_k2.log 2>&1
echo "rc=$?"
