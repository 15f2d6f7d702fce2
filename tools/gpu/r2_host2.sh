#!/bin/bash
# after a change of the host path: whole GPU suite, host profile, per-warp rasterizer timeline, split experiments
mkdir -p gpurun_out
tag=${1:-r2m2}
timeout 1800 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_tests.log 2>&1
tail -5 gpurun_out/${tag}_tests.log
timeout 300 python tools/host_prof.py C2 300 > gpurun_out/${tag}_host.txt 2>&1; head -4 gpurun_out/${tag}_host.txt; grep "per call" gpurun_out/${tag}_host.txt
timeout 300 python tools/trace_raster.py C2 > gpurun_out/${tag}_trace.txt 2>&1; head -24 gpurun_out/${tag}_trace.txt
echo "== FWD_SPLIT=2 trace"; GI2D_FWD_SPLIT=2 timeout 300 python tools/trace_raster.py C2 2>&1 | head -12
for v in "GI2D_FWD_SPLIT=2" "GI2D_BWD_SPLIT=4" "GI2D_FWD_CTAS_PER_SM=2"; do
  for c in C2 C3; do
    echo "== $v $c"; env $v timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "raster_"
  done
done
