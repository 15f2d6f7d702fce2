mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -5 gpurun_out/q_tests.log
for v in 1 0; do echo "== GI2D_FWD_WS=$v"
GI2D_FWD_WS=$v timeout 200 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster_fwd|total" | cut -c1-100
GI2D_FWD_WS=$v timeout 200 python tools/kernel_times.py C2 fit 20 2>&1 | grep -E "raster_fwd|total" | cut -c1-100
GI2D_FWD_WS=$v timeout 200 python tools/kernel_times.py C3 fit 20 2>&1 | grep -E "raster_fwd|total" | cut -c1-100
GI2D_FWD_WS=$v timeout 200 python tools/kernel_times.py C1 fit 20 2>&1 | grep -E "raster_fwd|total" | cut -c1-100
GI2D_FWD_WS=$v timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -E "raster_fwd|bin_small|total" | cut -c1-110
done
