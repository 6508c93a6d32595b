mkdir -p gpurun_out
timeout 240 python tools/tc_probe.py > gpurun_out/probe0.log 2>&1; echo "probe0 rc=$?"; tail -3 gpurun_out/probe0.log
timeout 600 python tools/tc_diag.py > gpurun_out/diag.log 2>&1; echo "diag rc=$?"; tail -60 gpurun_out/diag.log
timeout 900 python -m pytest tests/test_gpu_tc.py -q > gpurun_out/tc_tests.log 2>&1; echo "tc tests rc=$?"
tail -30 gpurun_out/tc_tests.log
