#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TMO=400 TAILN=25 run t_hs python -m pytest tests/test_gpu_headshard.py -m gpu -q -x --timeout 300
TMO=600 run t_all python -m pytest tests -m gpu -q --timeout 300
