#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_occ.py -q -m gpu -x > gpurun_out/occ_tests.log 2>&1; echo "pytest rc=$?"
grep -n "^E  " gpurun_out/occ_tests.log | head -30; tail -8 gpurun_out/occ_tests.log
