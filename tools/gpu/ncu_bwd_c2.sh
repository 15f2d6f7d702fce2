mkdir -p gpurun_out
timeout 200 python tools/prof_all.py c2 > gpurun_out/plain_prof_c2.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:raster_bwd -c 1 -f -o gpurun_out/r01_bwd_c2 python tools/prof_all.py c2 > gpurun_out/ncu_bwd_c2.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:bin_small -c 1 -f -o gpurun_out/r01_binsmall_c2 python tools/prof_all.py c2 > gpurun_out/ncu_binsmall_c2.log 2>&1
ls -la gpurun_out/*.ncu-rep
