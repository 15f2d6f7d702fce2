#!/bin/bash
# ncu evidence of the final kernels: launch list of the bench command, one-launch captures of the C2 chain, full sets for C2 and C5
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_launches.log 2>&1
bash tools/gpu/r2_ncu_raster.sh r02_ncu C2 > gpurun_out/r02_ncu_raster_stdout.txt 2>&1
bash tools/gpu/r2_ncu.sh r02full c5 > gpurun_out/r02_ncu_full_stdout.txt 2>&1
ls -la gpurun_out | grep "r02_ncu\|r02full" | awk '{print $5, $9}'
