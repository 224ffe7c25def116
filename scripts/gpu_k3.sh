#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TMO=300 run t_heat python -m pytest tests/test_gpu_heatmaps.py -m gpu -q -x --timeout 120
TMO=300 run t_engine python -m pytest tests/test_gpu_engine.py -m gpu -q --timeout 200
TMO=300 TAILN=20 run bench_heatmap python scripts/bench_heatmap.py
