mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_tests.log 2>&1; tail -8 gpurun_out/q_tests.log
timeout 300 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/q_decode8k.log 2>&1; tail -3 gpurun_out/q_decode8k.log | cut -c1-1500
