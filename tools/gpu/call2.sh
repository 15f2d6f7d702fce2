set -x; mkdir -p gpurun_out
python tests/golden/make_golden_refcuda.py gpurun_out/ref_cuda_kernels.npz 2>&1 | tail -3
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/s11b_tests.log 2>&1; tail -3 gpurun_out/s11b_tests.log
for c in C2 C1 C3; do timeout 300 python tests/bench_vs_refcuda.py --config $c --out gpurun_out/vs_refcuda_$c.json > gpurun_out/vs_refcuda_$c.log 2>&1; tail -12 gpurun_out/vs_refcuda_$c.log; done
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_ncu_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
timeout 200 python tools/prof_step.py C2 6 > gpurun_out/plain_prof.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'raster_bwd|raster_fwd|onesweep|scan_kernel|project_' -s 40 -c 8 -o gpurun_out/r01_full_C2 python tools/prof_step.py C2 6 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
