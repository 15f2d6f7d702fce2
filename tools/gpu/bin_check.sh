# after a change to the binning kernels: the parity tests that look at tile lists / bin edges / images, then kernel times
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5) > gpurun_out/bin_check.log 2>&1
for c in C2 C3 C1; do
  echo "=== $c" >> gpurun_out/bin_check.log
  (timeout 200 python tools/kernel_times.py $c render 20 2>&1 | grep "us/step" | head -4) >> gpurun_out/bin_check.log 2>&1
done
cat gpurun_out/bin_check.log
