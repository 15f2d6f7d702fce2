#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (its file is collected last)
mkdir -p gpurun_out
tag=${1:-r2n}
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/${tag}_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_tests.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/${tag}_tests.log
for c in C2 C3 C1; do
  echo "== $c"
  timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "raster_|bin_small|project|total"
done
echo "== C5 render"
timeout 300 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster_|total"
