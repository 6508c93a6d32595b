#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tc.py -q -m gpu -k "early_termination" > gpurun_out/et_tests.log 2>&1; echo "pytest rc=$?"; grep "^E  " gpurun_out/et_tests.log | head; tail -3 gpurun_out/et_tests.log
timeout 600 python tools/render_bench.py 2>&1 | tail -16
