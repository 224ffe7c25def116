#!/bin/bash
# K2 variants: 0 = v3 25% poly, 10 = v3 37.5% poly, 9 = v4 rolled MMA, 11 = v4 unrolled MMA, 12 = v4 unrolled 37.5%
mkdir -p gpurun_out
timeout 60 python scripts/bench_attention.py --impls 11 --iters 3 --only "cfg2 single" > gpurun_out/attn5_quick.log 2>&1; rc=$?; echo "quick rc=$rc"; cut -c1-260 gpurun_out/attn5_quick.log
[ $rc -ne 0 ] && exit 1
timeout 300 python scripts/bench_attention.py --impls 0,10,9,11,12 --iters 30 > gpurun_out/attn5.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/attn5.log'):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(d['impl'], d['shape'], d['ms'], d['TFLOPs'])
PY
