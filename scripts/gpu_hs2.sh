#!/bin/bash
# tensor-parallel head-sharded mode: 2-process parity test on one GPU, then config 5 at N=1 and N=$1
mkdir -p gpurun_out
N=${1:-2}
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-4} gpurun_out/$name.log | cut -c1-1800; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
TMO=300 TAILN=15 run t_hs python -m pytest tests/test_gpu_headshard.py -m gpu -q -x --timeout 200
TMO=400 run hs${N} $TR bench.py --gpus $N --config schnell-2048-hs --steps 2 --warmup 3 --no-cpu-baseline
