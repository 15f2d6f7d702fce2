mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -12 gpurun_out/q_tests.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/q_bench.log 2>&1; tail -2 gpurun_out/q_bench.log | cut -c1-3000
