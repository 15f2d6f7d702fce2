#!/bin/bash
# Round-2 evidence, one GPU: suite, smoke, both bench arms, kernel times, graph timeline, host profile, launch list, ncu captures.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r02_tests.log 2>&1; tail -4 gpurun_out/r02_tests.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/r02_smoke.log 2>&1; tail -2 gpurun_out/r02_smoke.log
timeout 400 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err
timeout 600 python bench.py --gpus 1 --steps 200 --warmup 20 --no-extras --no-cpu-baseline > gpurun_out/r02_bench_n1_k200.json 2> /dev/null
timeout 300 python bench.py --workload render --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_render_n1.json 2> /dev/null
timeout 300 python bench.py --workload decode8k --steps 30 --warmup 5 > gpurun_out/r02_decode8k_n1.json 2> /dev/null
timeout 400 python bench.py --workload kodak --steps 200 --warmup 20 > gpurun_out/r02_kodak_n1.json 2> /dev/null
for c in C1 C2 C3; do echo "== $c fit"; timeout 300 python tools/kernel_times.py $c fit 20 2>&1 | grep -E "us/step"; done > gpurun_out/r02_kernel_times.txt 2>&1
(echo "== C5 render"; timeout 300 python tools/kernel_times.py C5 render 10 2>&1 | grep -E "us/step") >> gpurun_out/r02_kernel_times.txt 2>&1
timeout 200 python tools/graph_timeline.py C2 > gpurun_out/r02_graph_timeline_C2.txt 2>&1
timeout 200 python tools/host_prof.py C2 300 > gpurun_out/r02_host_prof.txt 2>&1
for c in C1 C3; do timeout 300 python tests/bench_vs_refcuda.py --config $c --reps 100 --out gpurun_out/r02_vs_refcuda_$c.json > /dev/null 2>&1; done
cp gpurun_out/_bench_vs_refcuda_C2.json gpurun_out/r02_vs_refcuda_C2.json 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r02_ncu_launches.log 2>&1
bash tools/gpu/r2_ncu_raster.sh r02_ncu C2 > gpurun_out/r02_ncu_raster_stdout.txt 2>&1
bash tools/gpu/r2_ncu.sh r02full c5 > gpurun_out/r02_ncu_full_stdout.txt 2>&1
ls gpurun_out | grep r02 | head -50
