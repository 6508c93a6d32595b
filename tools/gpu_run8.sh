mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tc.py tests/test_gpu_engine.py tests/test_gpu_training.py -q -x > gpurun_out/t8.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/t8.log
for c in 1 2 4 8; do
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-render --bwd-chunks $c > gpurun_out/bench8_$c.json 2> gpurun_out/bench8_$c.err; echo "bench chunks=$c rc=$?"; tail -2 gpurun_out/bench8_$c.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench8_$c.json'))
print('chunks $c train', round(d['value']), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']))
PY
done
