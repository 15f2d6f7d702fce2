python tools/kernel_times.py C5 render 10 2>&1 | grep -E "raster|total" | cut -c1-100
python tools/kernel_times.py C2 fit 20 2>&1 | grep -E "raster|total" | cut -c1-100
python tools/kernel_times.py C3 fit 20 2>&1 | grep -E "raster|total" | cut -c1-100
python tools/kernel_times.py C1 fit 20 2>&1 | grep -E "raster|total" | cut -c1-100
timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -E "raster|total" | cut -c1-110
