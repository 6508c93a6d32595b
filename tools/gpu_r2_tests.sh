#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/t_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/t_tests.log
grep -n "^E  \|^FAILED\|passed\|failed\|rc=" gpurun_out/t_tests.log | head -60
tail -14 gpurun_out/t_tests.log
