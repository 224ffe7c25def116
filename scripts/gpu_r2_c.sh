#!/bin/bash
# round 2, call C: evidence for profiles/ — MMA/TMEM microbenchmark, K2 kernels vs SDPA, ncu of the default K2
mkdir -p gpurun_out
timeout 120 ./scripts/micro/umma_rate > gpurun_out/c_umma.md 2>&1; echo "umma rc=$?"
timeout 400 python scripts/bench_attention.py --sdpa --impls 1,2,3 > gpurun_out/c_k2_vs_sdpa.jsonl 2> gpurun_out/c_k2.err; echo "k2 rc=$?"
python scripts/bench_attention.py --only "cfg2 single" --impls 3 --iters 3 > /dev/null 2>&1 && \
timeout 400 ncu --set full --import-source on --clock-control none -k regex:attn_v11 -c 2 -o gpurun_out/c_k2v11 python scripts/bench_attention.py --only "cfg2 single" --impls 3 --iters 3 > gpurun_out/c_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/c_k2v11.ncu-rep --page source --csv > gpurun_out/c_src_v11.csv 2>/dev/null
ncu -i gpurun_out/c_k2v11.ncu-rep --page raw --csv > gpurun_out/c_raw_v11.csv 2>/dev/null
python scripts/bench_attention.py --only "cfg3" --impls 3 --iters 3 > /dev/null 2>&1 && \
timeout 400 ncu --set full --import-source on --clock-control none -k regex:attn_v11 -c 1 -o gpurun_out/c_k2v11_cfg3 python scripts/bench_attention.py --only "cfg3" --impls 3 --iters 3 > gpurun_out/c_ncu3.log 2>&1; echo "ncu3 rc=$?"
ncu -i gpurun_out/c_k2v11_cfg3.ncu-rep --page raw --csv > gpurun_out/c_raw_v11_cfg3.csv 2>/dev/null
