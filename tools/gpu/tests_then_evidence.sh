# round-end run: the whole GPU test suite first, then everything final_evidence.sh refreshes
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6) > gpurun_out/gpu_tests.log 2>&1; cat gpurun_out/gpu_tests.log
bash tools/gpu/final_evidence.sh
