mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_engine.py -q -x > gpurun_out/t5.log 2>&1; echo "tc+engine tests rc=$?"; tail -8 gpurun_out/t5.log
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench5.json 2> gpurun_out/bench5.err; echo "bench rc=$?"; tail -3 gpurun_out/bench5.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench5.json'))
print('train', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'render', d['render']['value'])
for k,v in d['kernels'].items(): print('  ',k, round(v['ms_per_step']*1e3,1),'us', v.get('frac'))
PY
