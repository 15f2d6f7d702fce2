# ncu --set full captures of the path's kernels at C2 (one process per kernel family); summaries go to profiles/.
set -x; mkdir -p gpurun_out
timeout 200 python tools/prof_step.py C2 6 > gpurun_out/plain_prof.log 2>&1 || exit 1
for k in raster_bwd raster_fwd scan_kernel onesweep project_fwd project_bwd; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 2 -f -o gpurun_out/r01_full_$k python tools/prof_step.py C2 6 > gpurun_out/ncu_full_$k.log 2>&1
  tail -2 gpurun_out/ncu_full_$k.log
done
