mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -6 gpurun_out/q_tests.log
timeout 300 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/q_decode8k.log 2>&1; tail -1 gpurun_out/q_decode8k.log | cut -c1-1800
