#!/bin/bash
# N-GPU check of the driver's scaling path: DP bench line and tensor-parallel config 5
mkdir -p gpurun_out
N=${1:-8}
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-1} gpurun_out/$name.log | cut -c1-1200; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
TMO=400 run dp${N} $TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline
TMO=400 run hs${N} $TR bench.py --gpus $N --config schnell-2048-hs --steps 2 --warmup 3 --no-cpu-baseline
