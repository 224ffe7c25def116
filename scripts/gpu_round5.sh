#!/bin/bash
# CUDA-graph path: engine tests, bench with and without graphs
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-6} gpurun_out/$name.log | cut -c1-${CUT:-400}; }
TMO=900 TAILN=25 run t_engine python -m pytest tests/test_gpu_engine.py tests/test_gpu_headshard.py -m gpu -q --timeout 400 -x
TMO=900 CUT=3000 TAILN=3 run bench_graph python bench.py --steps 3 --warmup 3 --no-cpu-baseline
TMO=900 CUT=3000 TAILN=1 run bench_nograph python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-graphs
TMO=600 CUT=3000 TAILN=3 run bench_enc python bench.py --config encode-512 --steps 4 --warmup 3 --no-cpu-baseline
