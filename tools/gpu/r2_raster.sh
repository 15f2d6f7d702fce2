#!/bin/bash
# After a change to the rasterizers: parity suite, then per-kernel times (CUDA events, L2 flushed) at C1/C2/C3/C5.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "not cpp_autograd_functions and not ctypes_host_path and not every_binning_path" > gpurun_out/r2b_tests.log 2>&1
tail -15 gpurun_out/r2b_tests.log
for c in C2 C1 C3; do timeout 300 python tools/kernel_times.py $c fit 20 > gpurun_out/r2b_kt_$c.txt 2>&1; tail -12 gpurun_out/r2b_kt_$c.txt; done
timeout 300 python tools/kernel_times.py C5 render 10 > gpurun_out/r2b_kt_C5.txt 2>&1; tail -8 gpurun_out/r2b_kt_C5.txt
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
python - <<'P'
import json
try:
    b=json.load(open('gpurun_out/r2b_bench.json')); print(b['value'], b['ms_per_step'], {k:round(v['us'],1) for k,v in b['stages'].items()}, b['extras'])
except Exception as e: print('bench failed', e); print(open('gpurun_out/r2b_bench.err').read()[-1500:])
P
