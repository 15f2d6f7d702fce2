#!/bin/bash
# Round 2, first call: the round-1 kernels under the NEW tests (full-size parity, caller-shaped model, C++ autograd suite),
# one bench line, the reference-CUDA comparison and the host profile on the same box -- the baseline every later change
# of the round is compared with.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/r2a_gpu.txt 2>&1
nproc >> gpurun_out/r2a_gpu.txt
timeout 1200 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r2a_tests.log 2>&1
tail -5 gpurun_out/r2a_tests.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
timeout 300 python tests/bench_vs_refcuda.py --config C2 --reps 200 --out gpurun_out/r2a_vs_refcuda_C2.json > /dev/null 2> gpurun_out/r2a_vs_refcuda.err
timeout 200 python tools/host_prof.py C2 300 > gpurun_out/r2a_host_prof.txt 2>&1
GI2D_CPP_AUTOGRAD=1 timeout 200 python tools/host_prof.py C2 300 > gpurun_out/r2a_host_prof_cpp.txt 2>&1
echo done
