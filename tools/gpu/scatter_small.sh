# does the scatter binning (default above 4096 tiles) pay below that?  C1 / C2 / C3 kernel times with the one-kernel path
# switched off and the scatter path forced on
mkdir -p gpurun_out
for c in C2 C3 C1; do
  echo "=== $c GI2D_SMALL_BIN=0 GI2D_SCATTER_BIN=2" >> gpurun_out/scatter_small.log
  (GI2D_SMALL_BIN=0 GI2D_SCATTER_BIN=2 timeout 200 python tools/kernel_times.py $c render 20 2>&1 | grep "us/step" | head -9) >> gpurun_out/scatter_small.log 2>&1
done
cat gpurun_out/scatter_small.log
