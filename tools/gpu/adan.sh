mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_tests.log 2>&1; tail -8 gpurun_out/q_tests.log
timeout 300 python tools/bench_adan.py --out gpurun_out/adan.json > gpurun_out/adan.log 2>&1; tail -30 gpurun_out/adan.log
