#!/bin/bash
# graph memset folding: its test, the graph timeline with and without it, one bench line each
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fit.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4
for f in 0 1; do
  echo "== GI2D_FOLD_MEMSETS=$f"
  GI2D_FOLD_MEMSETS=$f timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -v "Warn\|warn" | cut -c1-100 | tail -14
  GI2D_FOLD_MEMSETS=$f timeout 300 python bench.py --steps 50 --warmup 5 --no-extras --no-cpu-baseline > gpurun_out/r2fold_$f.json 2> gpurun_out/r2fold_$f.err
  python - <<P
import json
try:
    b=json.loads(open('gpurun_out/r2fold_$f.json').read().strip().splitlines()[-1]); print('value', b['value'], 'ms', b['ms_per_step'], b['ms_per_step_spread'], 'e2e', b['e2e']['value'], b.get('graph'))
except Exception as e: print('bench failed', e); print(open('gpurun_out/r2fold_$f.err').read()[-1500:])
P
done
