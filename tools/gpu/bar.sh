mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -4 gpurun_out/q_tests.log
for i in 1 2 3 4 5; do for v in 1 0; do echo -n "own=$v  "; GI2D_OWN_BARRIER=$v timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -E "bin_small" | cut -c1-60; done; done
