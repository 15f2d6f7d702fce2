#!/bin/bash
# sharded 8K decode at N ranks (standalone workload line) + parity of the row-sharded tests.  bash tools/gpu/r2_decode.sh <N> <tag>
N=${1:-2}; tag=${2:-r2d8}
mkdir -p gpurun_out
if [ "$N" == "1" ]; then
  timeout 600 python bench.py --workload decode8k --steps 30 --warmup 5 > gpurun_out/${tag}_n$N.json 2> gpurun_out/${tag}_n$N.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --workload decode8k --gpus $N --steps 30 --warmup 5 > gpurun_out/${tag}_n$N.json 2> gpurun_out/${tag}_n$N.err
fi
grep -v "Warning\|warn\|OMP_NUM\|\*\*\*\*" gpurun_out/${tag}_n$N.err | tail -12
python - <<P
import json
try:
    b=json.loads(open('gpurun_out/${tag}_n$N.json').read().strip().splitlines()[-1])
    print('value', b['value'], 'ms', b['ms_per_step'], 'e2e', b['e2e']['value']); print(b['summary']); print({k:v for k,v in b['config'].items() if 'ms' in k})
except Exception as e: print('bench failed', e)
P
