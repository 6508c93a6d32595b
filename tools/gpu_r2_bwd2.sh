#!/bin/bash
# tensor-core backward: parity tests (one tile in flight = default, then two), then the A/B timing with its probes
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 200 python -m pytest -x -q -m gpu tests/test_gpu_tc.py tests/test_gpu_training.py > gpurun_out/w_tc1.log 2>&1; echo "tc tests (default) rc=$?"; tail -3 gpurun_out/w_tc1.log
ACN_BWD_TILES=2 timeout 200 python -m pytest -x -q -m gpu tests/test_gpu_tc.py -k "tc_single or tc_eight or ragged or deterministic" > gpurun_out/w_tc2.log 2>&1; echo "tc tests (2 tiles) rc=$?"; tail -3 gpurun_out/w_tc2.log
timeout 400 python tools/bwd_bench.py 2>&1 | grep -v "^W\|Warning" | tee gpurun_out/w_bwd_bench.txt
