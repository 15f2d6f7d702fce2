set -x; mkdir -p gpurun_out
timeout 200 python tools/prof_step.py C2 6 render > gpurun_out/plain_prof.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:bin_small -s 3 -c 1 -f -o gpurun_out/r01_full_bin_small python tools/prof_step.py C2 6 render > gpurun_out/ncu_bs.log 2>&1
tail -2 gpurun_out/ncu_bs.log
