#!/bin/bash
# round 2, call B: the whole GPU suite + bench with the new K2 default
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -s > gpurun_out/b_pytest.log 2>&1
echo "pytest rc=$?" > gpurun_out/b_host.txt
tail -8 gpurun_out/b_pytest.log
timeout 700 python bench.py > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err
echo "bench rc=$?" >> gpurun_out/b_host.txt
tail -c 2500 gpurun_out/b_bench.json
cat gpurun_out/b_host.txt
