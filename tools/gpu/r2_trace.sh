#!/bin/bash
# after a rasterizer change: rasterizer parity tests, per-warp timeline, kernel times
mkdir -p gpurun_out
tag=${1:-r2t}
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "not cpp_autograd_functions and not ctypes_host_path and not every_binning_path" > gpurun_out/${tag}_tests.log 2>&1
tail -3 gpurun_out/${tag}_tests.log
timeout 300 python tools/trace_raster.py ${2:-C2} > gpurun_out/${tag}_trace.txt 2>&1; grep -A4 "^== " gpurun_out/${tag}_trace.txt | head -24
echo "== C5 render"; timeout 300 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster_"
for c in C2 C3 C1; do
  echo "== $c"; timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "raster_"
done
