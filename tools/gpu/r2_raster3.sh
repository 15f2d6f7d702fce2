#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider -k "not cpp_autograd_functions and not ctypes_host_path and not every_binning_path" > gpurun_out/r2f_tests.log 2>&1
tail -6 gpurun_out/r2f_tests.log
for v in "1" "2"; do
  for c in C2 C3 C1; do
    echo "== BWD_SPLIT=$v $c"
    GI2D_FWD_SPLIT=1 GI2D_BWD_SPLIT=$v timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "raster_|total"
  done
done
