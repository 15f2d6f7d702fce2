# A/B of library variants built into gaussianimage_b200/_variants/ (libgi2d_<name>.so): for each, the C2 graph timeline
# and the C3 fit kernel times.  usage: bash tools/gpu/variants.sh base b5 b6
mkdir -p gpurun_out
cp gaussianimage_b200/libgi2d.so /tmp/libgi2d_keep.so
for v in "$@"; do
  cp gaussianimage_b200/_variants/libgi2d_$v.so gaussianimage_b200/libgi2d.so
  echo "=== $v" | tee -a gpurun_out/variants.log
  (timeout 200 python tools/graph_timeline.py C2 2>&1 | tail -14) >> gpurun_out/variants.log 2>&1
  (timeout 200 python tools/kernel_times.py C3 fit 20 2>&1 | grep -i "raster\|bin_small\|total" | head -6) >> gpurun_out/variants.log 2>&1
  (timeout 200 python tools/kernel_times.py C1 fit 20 2>&1 | grep -i "raster\|total" | head -4) >> gpurun_out/variants.log 2>&1
done
cp /tmp/libgi2d_keep.so gaussianimage_b200/libgi2d.so
cat gpurun_out/variants.log
