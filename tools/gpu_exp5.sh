mkdir -p gpurun_out


timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_all.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b.json 2> gpurun_out/b.err || tail -3 gpurun_out/b.err
python - <<PY
import json
d=json.load(open('gpurun_out/b.json'))
print('train', round(d['value']), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), 'render', round(d['render']['value']), 'c5', round(d['render_c5']['value']), 'loss', d['final_loss'])
for k,v in d['kernels'].items(): print('  ',k, round(v['ms_per_step']*1e3,1),'us', v.get('frac'))
PY
