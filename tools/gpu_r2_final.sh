#!/bin/bash
# round-2 acceptance run on one GPU: smoke(), the whole GPU suite, the default bench line and the reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/f_smoke.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/f_tests.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/f_tests.log
timeout 900 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/f_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err; echo "reference arm rc=$?"; head -c 700 gpurun_out/f_ref.json
