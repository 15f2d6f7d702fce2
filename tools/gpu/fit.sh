mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_gpu_fit.py -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -12 gpurun_out/q_tests.log
timeout 600 python bench.py --workload kodak --steps 500 > gpurun_out/q_kodak.log 2>&1; tail -2 gpurun_out/q_kodak.log | cut -c1-2500
