mkdir -p gpurun_out
timeout 300 python bench.py --workload decode8k --no-cpu-baseline > gpurun_out/q_decode8k.log 2>&1; tail -1 gpurun_out/q_decode8k.log | cut -c1-300
timeout 600 python bench.py --workload kodak --steps 500 > gpurun_out/q_kodak.log 2>&1; tail -1 gpurun_out/q_kodak.log | cut -c1-300
python tools/kernel_times.py C5 render 10 2>&1 | grep -E "gi2d|total" | cut -c1-100
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
