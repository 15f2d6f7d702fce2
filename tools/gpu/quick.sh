# quick check after a kernel change: GPU tests, graph timeline (C2), decode breakdown, one bench line
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/q_tests.log 2>&1; tail -3 gpurun_out/q_tests.log
timeout 200 python tools/graph_timeline.py C2 > gpurun_out/q_timeline.log 2>&1; tail -16 gpurun_out/q_timeline.log
timeout 300 python tools/decode_breakdown.py > gpurun_out/q_decode.log 2>&1; tail -12 gpurun_out/q_decode.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/q_bench.log 2>&1; tail -1 gpurun_out/q_bench.log
