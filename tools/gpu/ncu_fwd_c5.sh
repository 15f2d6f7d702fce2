mkdir -p gpurun_out
timeout 200 python tools/prof_all.py c5 > gpurun_out/plain_prof_c5.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:raster_fwd -c 1 -f -o gpurun_out/r01_fwd_c5 python tools/prof_all.py c5 > gpurun_out/ncu_fwd_c5.log 2>&1
tail -2 gpurun_out/ncu_fwd_c5.log; ls -la gpurun_out/*.ncu-rep
