#!/bin/bash
# round 2, call E (N GPUs): config 5 sharded over N GPUs, exchange through peer memory vs NCCL all-to-all
mkdir -p gpurun_out
N=${1:-8}
for tag in peer nccl; do
  extra=""; [ $tag = nccl ] && extra="--no-peer-exchange"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N \
    bench.py --gpus $N --config schnell-2048-hs --steps 3 --warmup 3 --no-cpu-baseline $extra > gpurun_out/e_${tag}_$N.log 2>&1
  echo "rc=$? $tag N=$N"
  grep -o "\"value\": [0-9.]*, \"unit\": \"images/s\", \"n_gpus\"\|\"e2e\": {\"value\": [0-9.]*\|\"graph_replay\": {\"enabled\": [a-z]*, \"value\": [0-9.]*\|\"sm_mhz\": [0-9.]*\|\"ms_per_step\": [0-9.]*\|\"avg_launch_ms\": [0-9.]*" gpurun_out/e_${tag}_$N.log | tr "\n" " "; echo
  grep -i "error\|Traceback\|warn" gpurun_out/e_${tag}_$N.log | head -5
done
