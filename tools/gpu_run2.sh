mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py -q > gpurun_out/tc_tests.log 2>&1; echo "tc tests rc=$?"; tail -5 gpurun_out/tc_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_tc.json 2> gpurun_out/bench_tc.err; echo "bench tc rc=$?"; tail -3 gpurun_out/bench_tc.err; cat gpurun_out/bench_tc.json
timeout 1200 python -m pytest tests/test_gpu_parity.py -q -x > gpurun_out/parity.log 2>&1; echo "parity rc=$?"; tail -5 gpurun_out/parity.log
