#!/bin/bash
# ncu launch list of the timed region of bench.py + one full capture of the K2 kernel.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/plain.log; exit 1; }
tail -n 2 gpurun_out/plain.log | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 6000 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches.csv
ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_v1 -s 20 -c 2 -o gpurun_out/k2_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
