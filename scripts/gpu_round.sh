#!/bin/bash
# Runs on the GPU box under gpurun: tests in isolation (a hung kernel must not hide the
# other results), then smoke and a short bench.  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
run() { name=$1; shift; echo "=== $name"; timeout "$TMO" "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $name"; tail -n ${TAILN:-15} gpurun_out/$name.log; }
TMO=300
run t_heatmaps python -m pytest tests/test_gpu_heatmaps.py -m gpu -q -x --timeout 120
run t_k1 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "ln_modulate or gate_residual or qkv_prep" --timeout 120
TMO=240 run t_attn_small python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k "test_joint_attention and not full_size" --timeout 100
TMO=240 run t_attn_full python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "full_size or one_way" --timeout 200
TMO=300 run t_engine python -m pytest tests/test_gpu_engine.py -m gpu -q --timeout 200
TMO=200 run smoke python __graft_entry__.py smoke
