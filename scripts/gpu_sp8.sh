#!/bin/bash
# config 5 on N GPUs, one or both head-shard modes: scripts/gpu_sp8.sh N "sequence tensor"
mkdir -p gpurun_out
N=${1:-8}
for mode in ${2:-sequence}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N \
    bench.py --gpus $N --config schnell-2048-hs --steps 3 --warmup 3 --no-cpu-baseline --hs-mode $mode > gpurun_out/sp_${mode}_$N.log 2>&1
  echo "rc=$? $mode N=$N"
  grep -o "\"value\": [0-9.]*, \"unit\": \"images/s\", \"n_gpus\"\|\"e2e\": {\"value\": [0-9.]*\|\"graph_replay\": {\"enabled\": [a-z]*, \"value\": [0-9.]*\|\"sm_mhz\": [0-9.]*\|\"ms_per_step\": [0-9.]*\|\"avg_launch_ms\": [0-9.]*" gpurun_out/sp_${mode}_$N.log | tr "\n" " "; echo
  grep -i "error\|Traceback" gpurun_out/sp_${mode}_$N.log | head -5
done
