mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_tc.py -q > gpurun_out/t3.log 2>&1; echo "engine+tc tests rc=$?"; tail -25 gpurun_out/t3.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench graph rc=$?"; tail -5 gpurun_out/bench_graph.err; cat gpurun_out/bench_graph.json
timeout 900 python bench.py --steps 20 --warmup 3 --no-graph --no-cpu-baseline --no-render > gpurun_out/bench_eager.json 2> gpurun_out/bench_eager.err; echo "bench eager rc=$?"; tail -5 gpurun_out/bench_eager.err; python -c "
import json;d=json.load(open('gpurun_out/bench_eager.json'));print('eager',d['value'],d['ms_per_step'],d['e2e']['value'])"
timeout 1200 python -m pytest tests/test_gpu_parity.py -q -x > gpurun_out/parity.log 2>&1; echo "parity rc=$?"; tail -5 gpurun_out/parity.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
