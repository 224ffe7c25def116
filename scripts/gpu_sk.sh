#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 100 -k "attention" > gpurun_out/t_sk.log 2>&1; rc=$?; echo "tests rc=$rc"; tail -5 gpurun_out/t_sk.log | cut -c1-300
[ $rc -ne 0 ] && exit 1
for sk in 0 -1; do
CA_STREAMK=$sk timeout 200 python scripts/bench_attention.py --impls 0 --iters 30 > gpurun_out/sk_bench_$sk.log 2>&1; echo "bench streamk=$sk rc=$?"
python - <<PY
import json
for l in open('gpurun_out/sk_bench_$sk.log'):
    try: d=json.loads(l)
    except Exception: continue
    print(d['shape'], d['ms'], d['TFLOPs'])
PY
done
