mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -5 gpurun_out/q_tests.log
python tools/kernel_times.py C5 render 10 2>&1 | grep -E "gi2d|total" | cut -c1-100
python tools/kernel_times.py C2 fit 20 2>&1 | grep -E "gi2d|total" | cut -c1-100
python tools/kernel_times.py C3 fit 20 2>&1 | grep -E "raster|total" | cut -c1-100
python tools/kernel_times.py C1 fit 20 2>&1 | grep -E "raster|total" | cut -c1-100
timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -E "gi2d|total" | cut -c1-110
