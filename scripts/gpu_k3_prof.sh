#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/bench_heatmap.py --impls 0 --splits 1 --only cfg5"
$CMD > gpurun_out/k3_plain.log 2>&1 || { tail gpurun_out/k3_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:heatmap_tc -s 10 -c 2 -o gpurun_out/k3_cfg5 -f $CMD > gpurun_out/ncu_k3.log 2>&1
echo "rc=$?"
CMD="python scripts/bench_heatmap.py --impls 0 --splits 1 --only cfg3"
$CMD > gpurun_out/k3_plain3.log 2>&1 || { tail gpurun_out/k3_plain3.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:heatmap_tc -s 10 -c 1 -o gpurun_out/k3_cfg3 -f $CMD > gpurun_out/ncu_k3b.log 2>&1
echo "rc=$?"
