mkdir -p gpurun_out
for combo in "4 4 1024" "8 8 1024" "16 8 1024" "16 16 1024" "16 16 512" "8 16 1024"; do
  set -- $combo
  echo "== PASS_ITEMS=$1 LAST_ITEMS=$2 SCAN_T=$3"
  GI2D_PASS_ITEMS=$1 GI2D_LAST_ITEMS=$2 GI2D_SCAN_T=$3 timeout 200 python tools/bench_chain.py C5 2>&1 | tail -8
done > gpurun_out/c5_items.log 2>&1
cat gpurun_out/c5_items.log
