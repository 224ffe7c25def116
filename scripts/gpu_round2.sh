#!/bin/bash
# tests (whole gpu suite), smoke, bench; logs -> gpurun_out/
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TMO=600 run t_all python -m pytest tests -m gpu -q --timeout 300
TMO=300 run smoke python __graft_entry__.py smoke
TMO=900 TAILN=5 run bench python bench.py --steps 3 --warmup 3
