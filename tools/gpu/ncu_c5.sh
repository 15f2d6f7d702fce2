set -x; mkdir -p gpurun_out
timeout 200 python tools/prof_step.py C5 3 render > gpurun_out/plain_prof_c5.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'raster_fwd|scan_kernel|onesweep' -s 4 -c 5 -f -o gpurun_out/r01_full_C5 python tools/prof_step.py C5 3 render > gpurun_out/ncu_full_c5.log 2>&1
tail -2 gpurun_out/ncu_full_c5.log
