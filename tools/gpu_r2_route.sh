#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python -m pytest -q -m gpu tests/test_gpu_parity.py tests/test_gpu_step.py tests/test_gpu_fullsize.py -x 2>&1 | tail -4
python tools/route_bench.py 20 2>&1 | tail -2 | tee gpurun_out/y_route_bench.txt
python tools/render_bench.py 2>&1 | tail -12 | tee gpurun_out/y_render_bench.txt
