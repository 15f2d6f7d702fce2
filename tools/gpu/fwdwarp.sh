mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25) > gpurun_out/q_tests.log 2>&1; tail -4 gpurun_out/q_tests.log
python tools/kernel_times.py C5 render 10 2>&1 | tail -9 | cut -c1-100
python tools/kernel_times.py C2 fit 20 2>&1 | tail -16 | cut -c1-100
python tools/kernel_times.py C3 fit 20 2>&1 | grep raster | cut -c1-100
python tools/kernel_times.py C1 fit 20 2>&1 | grep raster | cut -c1-100
echo "== old kernel"; GI2D_FWD_WARP=0 python tools/kernel_times.py C2 fit 20 2>&1 | grep raster_fwd | cut -c1-100
GI2D_FWD_WARP=0 python tools/kernel_times.py C3 fit 20 2>&1 | grep raster_fwd | cut -c1-100
GI2D_FWD_WARP=0 python tools/kernel_times.py C1 fit 20 2>&1 | grep raster_fwd | cut -c1-100
