mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -12 gpurun_out/q_tests.log
timeout 200 python tools/bench_chain.py C5 2>&1 | tail -5
GI2D_SMALL_BIN=0 GI2D_SCATTER_BIN=2 timeout 200 python tools/bench_chain.py C2 2>&1 | tail -5
timeout 300 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/q_decode8k.log 2>&1; tail -1 gpurun_out/q_decode8k.log | cut -c1-400
