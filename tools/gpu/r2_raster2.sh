#!/bin/bash
# parity (quick selection) + kernel-time matrix over the forward / backward warp-split variants
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "not cpp_autograd_functions and not ctypes_host_path and not every_binning_path" > gpurun_out/r2d_tests.log 2>&1
tail -6 gpurun_out/r2d_tests.log
for v in "1 1" "2 2" "2 4" "2 1"; do
  set -- $v
  for c in C2 C3 C1; do
    echo "== FWD_SPLIT=$1 BWD_SPLIT=$2 $c"
    GI2D_FWD_SPLIT=$1 GI2D_BWD_SPLIT=$2 timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "raster_|total"
  done
done
echo "== defaults C5"
timeout 300 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster_|total"
GI2D_FWD_SPLIT=2 timeout 300 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster_|total"
