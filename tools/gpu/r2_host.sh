#!/bin/bash
# After a change to the host path (csrc/torch_glue.cpp, rasterize_sum.py): whole GPU suite, host profile, one bench line.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r2h_tests.log 2>&1
tail -25 gpurun_out/r2h_tests.log
timeout 200 python tools/host_prof.py C2 300 > gpurun_out/r2h_host_prof.txt 2>&1; head -5 gpurun_out/r2h_host_prof.txt; sed -n 6,30p gpurun_out/r2h_host_prof.txt | cut -c1-150
GI2D_CPP_AUTOGRAD=0 timeout 200 python tools/host_prof.py C2 300 2>&1 | head -4
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err
python - <<'P'
import json
try:
    b=json.load(open('gpurun_out/r2h_bench.json')); print(b['value'], b['ms_per_step'], b['e2e'], {k:round(v['us'],1) for k,v in b['stages'].items()}, b['extras'])
except Exception as e: print('bench failed', e); print(open('gpurun_out/r2h_bench.err').read()[-2500:])
P
