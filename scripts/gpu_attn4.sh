#!/bin/bash
# K2 v4 vs v3: quick hang check, kernel parity tests, then the micro-benchmark (impl 0/6/7 = v4, 8/4/5 = v3)
mkdir -p gpurun_out
timeout 90 python scripts/bench_attention.py --impls 0 --iters 3 --only "cfg2 single" > gpurun_out/attn4_quick.log 2>&1; rc=$?; echo "quick rc=$rc"; cut -c1-260 gpurun_out/attn4_quick.log
[ $rc -ne 0 ] && exit 1
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --timeout 120 > gpurun_out/t_attn.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_attn.log
timeout 300 python scripts/bench_attention.py --impls 0,6,7,8,4 --iters 30 > gpurun_out/attn4.log 2>&1; echo "bench rc=$?"
cat gpurun_out/attn4.log | cut -c1-260
