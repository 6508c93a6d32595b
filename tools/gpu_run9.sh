mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_all.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench9.json 2> gpurun_out/bench9.err; echo "bench rc=$?"; tail -3 gpurun_out/bench9.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench9.json'))
print('train', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'render', d['render'], 'c5', d['render_c5'])
for k,v in d['kernels'].items(): print('  ',k, round(v['ms_per_step']*1e3,1),'us', v.get('frac'))
print(d['roofline'], d['cpu_baseline'])
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400
