#!/bin/bash
# ncu --set full of every kernel of one C2 fit iteration (tools/prof_all.py c2) and of the C5 decode, exported to CSV on the box;
# plus the source-level hot spots of the two rasterizers.  Usage: bash tools/gpu/r2_ncu.sh <tag>
tag=${1:-r2}
mkdir -p gpurun_out
timeout 120 python tools/prof_all.py c2 > gpurun_out/${tag}_plain_c2.log 2>&1 || { tail -5 gpurun_out/${tag}_plain_c2.log; exit 1; }
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o gpurun_out/${tag}_full_c2 python tools/prof_all.py c2 > gpurun_out/${tag}_ncu_c2.log 2>&1
ncu -i gpurun_out/${tag}_full_c2.ncu-rep --page raw --csv > gpurun_out/${tag}_full_c2.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${tag}_full_c2.csv > gpurun_out/${tag}_ncu_c2.md 2>/dev/null
cat gpurun_out/${tag}_ncu_c2.md | cut -c1-400
for k in raster_fwd_kernel raster_bwd_kernel bin_small_kernel; do
  timeout 120 python tools/ncu_source_hot.py gpurun_out/${tag}_full_c2.ncu-rep $k > gpurun_out/${tag}_hot_$k.txt 2>&1
done
rm -f gpurun_out/${tag}_full_c2.ncu-rep
if [ "$2" == "c5" ]; then
timeout 400 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o gpurun_out/${tag}_full_c5 python tools/prof_all.py c5 > gpurun_out/${tag}_ncu_c5.log 2>&1
ncu -i gpurun_out/${tag}_full_c5.ncu-rep --page raw --csv > gpurun_out/${tag}_full_c5.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${tag}_full_c5.csv > gpurun_out/${tag}_ncu_c5.md 2>/dev/null
timeout 120 python tools/ncu_source_hot.py gpurun_out/${tag}_full_c5.ncu-rep raster_fwd_kernel > gpurun_out/${tag}_hot_c5_fwd.txt 2>&1
rm -f gpurun_out/${tag}_full_c5.ncu-rep
fi
