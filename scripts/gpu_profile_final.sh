#!/bin/bash
# ncu launch list of the timed region of bench.py (launch by launch region) + full capture of K2, current HEAD
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/plain.log; exit 1; }
tail -n 1 gpurun_out/plain.log | cut -c1-300
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 6000 \
    --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches_final.csv
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_v1 -s 20 -c 2 -o gpurun_out/k2_final -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:ln_stream -s 10 -c 2 -o gpurun_out/k1c_final -f $CMD > gpurun_out/ncu_k1c.log 2>&1
echo "ncu k1c rc=$?"; ls -la gpurun_out/*.ncu-rep
