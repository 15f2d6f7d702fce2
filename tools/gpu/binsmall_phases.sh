for d in 8 16 32 64 0; do echo "== GI2D_DBG=$d"; GI2D_DBG=$d timeout 200 python tools/bench_chain.py C2 2>&1 | grep binning; done
for nt in 256 512 1024; do echo "== NT=$nt"; GI2D_SMALL_NT=$nt timeout 200 python tools/bench_chain.py C2 2>&1 | grep binning; done
timeout 300 python tools/bench_adan.py 2>&1 | grep -E "fused|foreach"
