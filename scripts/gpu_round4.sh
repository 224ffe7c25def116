#!/bin/bash
# round-1 re-validation of HEAD: gpu tests, smoke, bench (both arms), ncu launch list, K2 full capture
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-6} gpurun_out/$name.log | cut -c1-${CUT:-400}; }
TMO=900 run t_all python -m pytest tests -m gpu -q --timeout 400
TMO=300 run smoke python __graft_entry__.py smoke
TMO=900 CUT=3000 TAILN=1 run bench python bench.py --steps 3 --warmup 3
TMO=600 CUT=3000 TAILN=1 run bench_ref python bench.py --impl reference --steps 1 --warmup 0
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 6000 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches.csv
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_v1 -s 20 -c 2 -o gpurun_out/k2_v3_full -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
