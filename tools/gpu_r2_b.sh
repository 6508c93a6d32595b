#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/adam_bench.py > gpurun_out/b_adam.txt 2>&1; cat gpurun_out/b_adam.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/b_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/b_tests.log
tail -15 gpurun_out/b_tests.log
