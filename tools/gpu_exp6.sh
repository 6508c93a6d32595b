mkdir -p gpurun_out
ACN_TC_DBG=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-render --no-graph > gpurun_out/dbg.json 2> gpurun_out/dbg.err
grep "TCDBG\|TCKERN" gpurun_out/dbg.err | tail -2
