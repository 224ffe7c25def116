#!/bin/bash
# round 2, call D (2 GPUs): NCCL parity test on real GPUs + the bench line at N=2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nccl.py -q -s > gpurun_out/d_nccl.log 2>&1; echo "nccl rc=$?"
tail -4 gpurun_out/d_nccl.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/d_bench2.json 2> gpurun_out/d_bench2.err; echo "bench2 rc=$?"
tail -c 1800 gpurun_out/d_bench2.json
