# bench lines (fit, render) after a change to bench.py's e2e leg, then the ncu --set full capture of the C2 kernels
mkdir -p gpurun_out
timeout 90 python bench.py --workload render --no-cpu-baseline > gpurun_out/e3_bench_render.json 2> gpurun_out/e3_bench_render.err; tail -c 700 gpurun_out/e3_bench_render.err
timeout 120 python bench.py > gpurun_out/e3_bench_n1.json 2> gpurun_out/e3_bench_n1.err
python - <<'P'
import json
for f in ("e3_bench_render", "e3_bench_n1"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1]); print(f, d["value"], d["e2e"])
    except Exception as e:
        print(f, "no line:", e)
P
timeout 100 python tools/prof_all.py c2 > gpurun_out/plain_prof_c2.log 2>&1 && \
timeout 200 ncu --set full --clock-control none --profile-from-start off -f -o gpurun_out/r01_full_c2 python tools/prof_all.py c2 > gpurun_out/ncu_full_c2.log 2>&1
ncu -i gpurun_out/r01_full_c2.ncu-rep --page raw --csv > gpurun_out/r01_full_c2.csv 2>/dev/null; rm -f gpurun_out/r01_full_c2.ncu-rep
tail -2 gpurun_out/ncu_full_c2.log
