mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_mlp_bwd_tc' -s 2 -c 1 -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu4.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu4.log
