mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/q_tests.log 2>&1; tail -6 gpurun_out/q_tests.log
for c in C2; do timeout 300 python tests/bench_vs_refcuda.py --config $c --out gpurun_out/vs_refcuda_$c.json > gpurun_out/vs_refcuda_$c.log 2>&1; tail -42 gpurun_out/vs_refcuda_$c.log; done
