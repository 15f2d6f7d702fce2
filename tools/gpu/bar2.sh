for i in 1 2 3; do echo -n "own=1  "; timeout 200 python tools/graph_timeline.py C2 2>&1 | grep -E "bin_small" | cut -c1-60; done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
