#!/bin/bash
# the driver's multi-GPU command at N ranks.  bash tools/gpu/r2_scale.sh <N> <tag>
N=${1:-2}; tag=${2:-r2s}
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${tag}_n$N.json 2> gpurun_out/${tag}_n$N.err ) 2> gpurun_out/${tag}_n$N.time
tail -3 gpurun_out/${tag}_n$N.time; grep -v "Warning\|warn\|return Variable" gpurun_out/${tag}_n$N.err | tail -15
python - <<P
import json
try:
    b=json.loads(open('gpurun_out/${tag}_n$N.json').read().strip().splitlines()[-1])
    print('value', b['value'], 'ms', b['ms_per_step'], 'e2e', b['e2e']['value'])
    x=b['extras']; print({k:x.get(k) for k in ('render_fps','kodak_imgs_per_s','us_per_training_iteration','render_error','kodak_error')}); print('decode8k', x.get('decode8k'))
except Exception as e: print('bench failed', e)
P
