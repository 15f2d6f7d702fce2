# refresh every number and profile kept under profiles/ (one B200)
set -x; mkdir -p gpurun_out
for c in C1 C2 C3; do timeout 300 python tests/bench_vs_refcuda.py --config $c --out gpurun_out/vs_refcuda_$c.json > gpurun_out/vs_refcuda_$c.log 2>&1; tail -3 gpurun_out/vs_refcuda_$c.log; done
timeout 300 python tools/bench_adan.py --out gpurun_out/adan.json > gpurun_out/adan.log 2>&1; tail -16 gpurun_out/adan.log
timeout 300 python bench.py --workload render --no-cpu-baseline > gpurun_out/bench_render.log 2>&1; tail -1 gpurun_out/bench_render.log | cut -c1-300
timeout 300 python bench.py > gpurun_out/bench_fit_full.log 2>&1; tail -1 gpurun_out/bench_fit_full.log | cut -c1-300
timeout 200 python tools/graph_timeline.py C2 > gpurun_out/graph_timeline_C2.txt 2>&1; tail -14 gpurun_out/graph_timeline_C2.txt
timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r01_ncu_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
timeout 200 python tools/prof_all.py c2 > gpurun_out/plain_prof_c2.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -f -o gpurun_out/r01_full_c2 python tools/prof_all.py c2 > gpurun_out/ncu_full_c2.log 2>&1
ncu -i gpurun_out/r01_full_c2.ncu-rep --page raw --csv > gpurun_out/r01_full_c2.csv 2>/dev/null; rm -f gpurun_out/r01_full_c2.ncu-rep
tail -2 gpurun_out/ncu_full_c2.log
timeout 200 python tools/prof_all.py c5 > gpurun_out/plain_prof_c5.log 2>&1 && \
ncu --set full --clock-control none --profile-from-start off -f -o gpurun_out/r01_full_c5 python tools/prof_all.py c5 > gpurun_out/ncu_full_c5.log 2>&1
ncu -i gpurun_out/r01_full_c5.ncu-rep --page raw --csv > gpurun_out/r01_full_c5.csv 2>/dev/null; rm -f gpurun_out/r01_full_c5.ncu-rep
tail -2 gpurun_out/ncu_full_c5.log
du -sh gpurun_out; ls -la gpurun_out | tail -30
