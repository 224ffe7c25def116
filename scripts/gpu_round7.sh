#!/bin/bash
# full gpu suite + bench (config 2) + config 5 on one GPU after the stream-K tail
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-6} gpurun_out/$name.log | cut -c1-${CUT:-400}; }
TMO=900 run t_all python -m pytest tests -m gpu -q --timeout 400
TMO=300 TAILN=2 run smoke python __graft_entry__.py smoke
TMO=900 CUT=3000 TAILN=1 run bench python bench.py --steps 3 --warmup 3 --no-cpu-baseline
