#!/bin/bash
# The driver's commands: both bench arms at N=1 (and the ncu evidence the line quotes).  bash tools/gpu/r2_bench.sh <tag>
tag=${1:-r2i}
mkdir -p gpurun_out
( time timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err ) 2> gpurun_out/${tag}_bench.time
tail -3 gpurun_out/${tag}_bench.time; tail -c 1500 gpurun_out/${tag}_bench.err
python - <<P
import json
try:
    b=json.load(open('gpurun_out/${tag}_bench.json'))
    print('value', b['value'], 'ms', b['ms_per_step'], b['ms_per_step_spread'])
    print('e2e', {k:v for k,v in b['e2e'].items() if k!='mode'})
    print('roofline', {k:v for k,v in b['roofline'].items() if k not in ('note','peak_source')})
    print('stages', {k:(round(v['us'],1), round(v['frac'],3)) for k,v in b['stages'].items()})
    print('extras', json.dumps(b['extras'], indent=0)[:3500])
    print('cpu', b['cpu_baseline'])
except Exception as e: print('bench failed', e)
P
