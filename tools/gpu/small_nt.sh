# block shape of the one-kernel binning: default (one block per SM) against the next smaller shape (two blocks per SM)
mkdir -p gpurun_out
run() { echo "=== $1 GI2D_SMALL_NT=$2" >> gpurun_out/small_nt.log
  (GI2D_SMALL_NT=$2 timeout 200 python tools/kernel_times.py $1 render 20 2>&1 | grep "bin_small" | head -3) >> gpurun_out/small_nt.log 2>&1; }
run C2 0; run C2 256; run C2 1024; run C3 0; run C3 512; run C3 256; run C1 0; run C1 512
cat gpurun_out/small_nt.log
