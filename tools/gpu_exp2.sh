mkdir -p gpurun_out
for c in 2 4; do
ACN_DEBUG_SERIAL_PIPE=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-render --bwd-chunks $c > gpurun_out/exp2_$c.json 2>/dev/null
python - <<PY
import json
d=json.load(open('gpurun_out/exp2_$c.json'))
print('SERIAL chunks $c train', round(d['value']), 'ms', round(d['ms_per_step'],4))
PY
done
