#!/bin/bash
# cumulative time of the one-kernel binning's phases (GI2D_DBG stops the kernel after a phase) at C2 / C3 / C1
for c in C2 C3 C1; do
  for d in 8 16 64 32 0; do
    echo -n "$c dbg=$d: "; GI2D_DBG=$d timeout 120 python tools/bench_chain.py $c 2>&1 | grep binning_chain
  done
done
