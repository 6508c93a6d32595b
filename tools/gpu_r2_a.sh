#!/bin/bash
# round 2, run A: GPU tests + first bench of the restructured headline
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/a_env.txt 2>&1
nproc >> gpurun_out/a_env.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_tests.log
tail -5 gpurun_out/a_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench rc=$?"
tail -c 1500 gpurun_out/a_bench.err
head -c 3000 gpurun_out/a_bench.json
