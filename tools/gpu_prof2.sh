mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_mlp' -s 4 -c 2 -o gpurun_out/prof_tc $CMD > gpurun_out/ncu3.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu3.log
