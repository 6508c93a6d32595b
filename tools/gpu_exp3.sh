mkdir -p gpurun_out
for v in 0; do
  ACN_HASH_BWD=$v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/hb_$v.json 2> gpurun_out/hb_$v.err || tail -3 gpurun_out/hb_$v.err
  python - <<PY
import json
d=json.load(open('gpurun_out/hb_$v.json'))
print('variant $v train', round(d['value']), 'ms', round(d['ms_per_step'],4), 'hash_bwd us', round(d['kernels']['hash_bwd']['ms_per_step']*1e3,1), 'loss', d['final_loss'])
PY
done
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_all.log
