# last call of a round on a short GPU budget: whole GPU suite, smoke(), both bench arms, launch list -- each step
# under its own timeout and writing its own file, most important first, so a cut-off call still leaves evidence
mkdir -p gpurun_out
(timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/re_tests.log 2>&1; cat gpurun_out/re_tests.log
timeout 150 python bench.py > gpurun_out/re_bench_n1.json 2> gpurun_out/re_bench_n1.err; tail -c 600 gpurun_out/re_bench_n1.json
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/re_smoke.log 2>&1; tail -1 gpurun_out/re_smoke.log
timeout 120 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/re_bench_ref.json 2> gpurun_out/re_bench_ref.err; tail -c 400 gpurun_out/re_bench_ref.json
timeout 90 python bench.py --workload render --no-cpu-baseline > gpurun_out/re_bench_render.json 2>/dev/null; tail -c 300 gpurun_out/re_bench_render.json
timeout 120 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/re_decode8k.json 2>/dev/null; tail -c 300 gpurun_out/re_decode8k.json
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/re_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/re_ncu.log 2>&1; tail -2 gpurun_out/re_ncu.log
