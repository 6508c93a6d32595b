#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it runs last)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/z_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/z_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/z_all.log 2>&1; echo "all rc=$?"; tail -5 gpurun_out/z_all.log
