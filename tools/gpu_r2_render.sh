#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/render_bench.py > gpurun_out/render_bench.txt 2>&1; echo "rc=$?"; tail -20 gpurun_out/render_bench.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_tc.py -q -m gpu -k "render or frame or 1080p or sample_major or fused or ragged" > gpurun_out/render_tests.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/render_tests.log
