#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TMO=300 run t_attn python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "attention or one_way" --timeout 120
TMO=300 run t_engine python -m pytest tests/test_gpu_engine.py tests/test_gpu_headshard.py -m gpu -q --timeout 200
TMO=300 TAILN=30 run bench_attn python scripts/bench_attention.py --impls 0,4,5,3
TMO=600 TAILN=2 run bench python bench.py --steps 3 --warmup 3 --no-cpu-baseline
