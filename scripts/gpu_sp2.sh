#!/bin/bash
# sequence-parallel mode: kernel + 2-process engine tests on one GPU, then config 5 on the GPUs given
mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m pytest tests/test_gpu_seqpar.py -x -q -s 2>&1 | tail -25
for mode in sequence tensor; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N \
    bench.py --gpus $N --config schnell-2048-hs --steps 3 --warmup 3 --no-cpu-baseline --hs-mode $mode > gpurun_out/sp_${mode}_$N.log 2>&1
  echo "rc=$? $mode N=$N"
  grep -o "\"value\": [0-9.]*, \"unit\": \"images/s\", \"n_gpus\"\|\"e2e\": {\"value\": [0-9.]*\|\"graph_replay\": {\"enabled\": [a-z]*, \"value\": [0-9.]*\|\"sm_mhz\": [0-9.]*\|\"ms_per_step\": [0-9.]*\|\"avg_launch_ms\": [0-9.]*" gpurun_out/sp_${mode}_$N.log | tr "\n" " "; echo
  grep -i "error\|Traceback" gpurun_out/sp_${mode}_$N.log | head -5
  tail -c 400 gpurun_out/sp_${mode}_$N.log | grep -v "^{" | tail -5
done
