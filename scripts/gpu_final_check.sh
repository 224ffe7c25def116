#!/bin/bash
# what the driver runs at round end: gpu tests, smoke, bench (default flags), reference arm
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-3} gpurun_out/$name.log | cut -c1-${CUT:-600}; }
TMO=900 run t_all python -m pytest tests -x -q -m gpu
TMO=300 TAILN=1 run smoke python -c "import __graft_entry__ as g; g.smoke()"
TMO=900 CUT=4000 TAILN=1 run bench_default python bench.py
TMO=600 CUT=1500 TAILN=1 run bench_ref python bench.py --impl reference --steps 5 --warmup 3
