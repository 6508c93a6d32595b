mkdir -p gpurun_out
for t in 0 128 256 384; do
ACN_TC_DBG=$t timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-render --no-graph > gpurun_out/dbg.json 2> gpurun_out/dbg.err
echo "tid $t"; grep "TCDBG" gpurun_out/dbg.err | tail -1 | cut -d' ' -f1-12,44-52
done
