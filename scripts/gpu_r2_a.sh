#!/bin/bash
# round 2, call A: parity at full depth + bench with the new keys + K2 vs the library SDPA
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/a_host.txt; nproc >> gpurun_out/a_host.txt
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv >> gpurun_out/a_host.txt
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_host.txt
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench rc=$?" >> gpurun_out/a_host.txt
tail -c 1500 gpurun_out/a_bench.json
timeout 300 python scripts/bench_attention.py --sdpa --impls 0 > gpurun_out/a_k2_vs_sdpa.jsonl 2> gpurun_out/a_k2_vs_sdpa.err
echo "sdpa rc=$?" >> gpurun_out/a_host.txt
cat gpurun_out/a_host.txt
