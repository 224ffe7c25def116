#!/bin/bash
# round 2: ncu launch list of the timed region of bench.py + full capture of the default K2 kernel inside it
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-other-configs --no-gpu-reference"
timeout 600 $CMD > gpurun_out/p_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/p_plain.log; exit 1; }
tail -n 1 gpurun_out/p_plain.log | cut -c1-300
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 8000 \
    --csv --log-file gpurun_out/p_launches.csv $CMD > gpurun_out/p_ncu_launches.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/p_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_v11_kernel -s 20 -c 2 -o gpurun_out/p_k2 -f $CMD > gpurun_out/p_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/p_k2.ncu-rep --page raw --csv > gpurun_out/p_k2_raw.csv 2>/dev/null
ncu -i gpurun_out/p_k2.ncu-rep --page source --csv > gpurun_out/p_k2_src.csv 2>/dev/null
ls -la gpurun_out/p_*
