mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -4 gpurun_out/q_tests.log
python tools/kernel_times.py C5 render 10 2>&1 | tail -9
timeout 300 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/q_decode8k.log 2>&1; tail -1 gpurun_out/q_decode8k.log | cut -c1-300
