# after a change to the host glue (csrc/torch_glue.cpp): whole GPU suite through it, then host time per eager step with
# and without it on the same box, then the reference-CUDA comparison at C1 (the config where eager fit was behind)
mkdir -p gpurun_out
(timeout 120 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/glue_tests.log 2>&1; cat gpurun_out/glue_tests.log
(timeout 40 python tools/host_prof.py C2 300 2>&1 | head -4) > gpurun_out/glue_host_on.log; cat gpurun_out/glue_host_on.log
(GI2D_HOST_GLUE=0 timeout 40 python tools/host_prof.py C2 300 2>&1 | head -4) > gpurun_out/glue_host_off.log; cat gpurun_out/glue_host_off.log
timeout 40 python tests/bench_vs_refcuda.py --config C1 --out gpurun_out/glue_vs_refcuda_C1.json > gpurun_out/glue_vs_refcuda_C1.log 2>&1; tail -9 gpurun_out/glue_vs_refcuda_C1.log
