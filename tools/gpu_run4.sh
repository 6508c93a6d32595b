mkdir -p gpurun_out
timeout 600 python tools/hash_levels.py > gpurun_out/levels.log 2>&1; echo "levels rc=$?"; cat gpurun_out/levels.log | tail -20
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -8 gpurun_out/gpu_tests.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench4.json 2> gpurun_out/bench4.err; echo "bench rc=$?"; tail -3 gpurun_out/bench4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench4.json'))
print('train', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'render', d['render']['value'])
for k,v in d['kernels'].items(): print('  ',k, round(v['ms_per_step']*1e3,1),'us', v.get('frac'))
PY
