#!/bin/bash
mkdir -p gpurun_out
for sk in 0 -1; do
CA_STREAMK=$sk timeout 200 ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_active.max,sm__cycles_active.min,sm__cycles_elapsed.max,launch__grid_size,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:attn_v1 -s 2 -c 2 --csv --log-file gpurun_out/sk_ncu_$sk.csv python scripts/bench_attention.py --impls 0 --iters 3 --only "3 heads" > gpurun_out/sk_ncu_$sk.log 2>&1; echo "rc=$?"
grep -v "^==" gpurun_out/sk_ncu_$sk.csv | python -c "
import csv,sys
for r in csv.DictReader(sys.stdin):
    print(r['ID'], r['Metric Name'], r['Metric Value'], r['Metric Unit'])
"
done
