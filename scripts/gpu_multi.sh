#!/bin/bash
# 2-GPU checks: DP bench (weak scaling), encode sweep, head-sharded config.
mkdir -p gpurun_out
N=${1:-2}
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-4} gpurun_out/$name.log | cut -c1-1500; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
TMO=600 run dp${N} $TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline
TMO=600 run enc1 python bench.py --config encode-512 --steps 4 --warmup 3 --no-cpu-baseline
TMO=600 run enc${N} $TR bench.py --gpus $N --config encode-512 --steps 4 --warmup 3
TMO=900 run hs1 python bench.py --config schnell-2048-hs --steps 2 --warmup 3 --no-cpu-baseline
TMO=900 run hs${N} $TR bench.py --gpus $N --config schnell-2048-hs --steps 2 --warmup 3
