mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pe.py -q -x > gpurun_out/pe.log 2>&1; echo "pe tests rc=$?"; tail -30 gpurun_out/pe.log
