#!/bin/bash
mkdir -p gpurun_out
timeout 60 python scripts/bench_attention.py --impls 13 --iters 3 --only "cfg2 single" > gpurun_out/q.log 2>&1; rc=$?; echo "quick rc=$rc"; tail -1 gpurun_out/q.log | cut -c1-200
[ $rc -ne 0 ] && exit 1
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 100 -k "attention" > gpurun_out/t_sk.log 2>&1; rc=$?; echo "tests rc=$rc"; tail -5 gpurun_out/t_sk.log | cut -c1-300
timeout 300 python scripts/bench_attention.py --impls 0,13 --iters 30 > gpurun_out/opt_bench.log 2>&1; echo "bench rc=$?"
python - <<PY
import json
for l in open('gpurun_out/opt_bench.log'):
    try: d=json.loads(l)
    except Exception: continue
    print(d['impl'], d['shape'], d['ms'], d['TFLOPs'])
PY
