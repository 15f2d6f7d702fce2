#!/bin/bash
# last run of the round: whole GPU suite, smoke, both bench arms at N=1 (the driver's commands)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/r02_tests.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02_tests.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r02_smoke.log 2>&1; tail -1 gpurun_out/r02_smoke.log
( time timeout 400 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err ) 2>&1 | grep real
( time timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err ) 2>&1 | grep real
python - <<'P'
import json
def last(f): return json.loads(open(f).read().strip().splitlines()[-1])
b=last('gpurun_out/r02_bench_n1.json'); a=last('gpurun_out/r02_bench_reference_arm.json')
print('value', round(b['value']), 'e2e', round(b['e2e']['value']), 'eager', round(b['e2e']['eager_strict_iters_per_s']), 'ref arm', a['value'], 'same config', a['config']==b['config'])
print('roofline', {k:b['roofline'][k] for k in ('kernel','frac','achieved','launch_us','traffic')}, (b['roofline'].get('ncu') or {}).get('warp_instructions'))
print('extras', {k:b['extras'].get(k) for k in ('render_fps','kodak_imgs_per_s')}, b['extras']['decode8k']['fps'], b['extras']['reference_cuda']['speedup_vs_reference_cuda']['fit_eager_strict'])
P
