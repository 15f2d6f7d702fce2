#!/bin/bash
# ncu --set full + source of ONE launch of bin / raster fwd / raster bwd at a config.  bash tools/gpu/r2_ncu_raster.sh <tag> [C2]
tag=${1:-r2e}; cfg=${2:-C2}
mkdir -p gpurun_out
timeout 120 python tools/prof_raster.py $cfg > gpurun_out/${tag}_plain.log 2>&1 || { tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o gpurun_out/${tag}_rep python tools/prof_raster.py $cfg > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/${tag}_rep.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${tag}_raw.csv > gpurun_out/${tag}_summary.md 2>/dev/null
cut -c1-420 gpurun_out/${tag}_summary.md
for k in raster_fwd_kernel raster_bwd_kernel bin_small_kernel; do
  ncu -i gpurun_out/${tag}_rep.ncu-rep --page source --csv --kernel-name regex:$k > gpurun_out/${tag}_src_$k.csv 2>/dev/null
done
ncu -i gpurun_out/${tag}_rep.ncu-rep --page details --csv > gpurun_out/${tag}_details.csv 2>/dev/null
rm -f gpurun_out/${tag}_rep.ncu-rep
