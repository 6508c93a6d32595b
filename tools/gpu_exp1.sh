mkdir -p gpurun_out
for g in 592 296 148 74 37 18; do
  ACN_DEBUG_HASH_BWD_GRID=$g timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-render --no-graph > gpurun_out/exp_$g.json 2>/dev/null
  python -c "
import json;d=json.load(open('gpurun_out/exp_$g.json'));print('grid.x',$g,'CTAs',$g*16,'hash_bwd us',round(d['kernels']['hash_bwd']['ms_per_step']*1e3,1))"
done
