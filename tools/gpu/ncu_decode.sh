set -x; mkdir -p gpurun_out
timeout 250 python bench.py --workload decode8k --steps 3 --warmup 3 > gpurun_out/plain_decode.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 120 --csv --log-file gpurun_out/r01_ncu_launches_decode8k.csv python bench.py --workload decode8k --steps 3 --warmup 3 > gpurun_out/ncu_decode.log 2>&1
tail -2 gpurun_out/ncu_decode.log
