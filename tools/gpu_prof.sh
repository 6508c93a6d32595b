mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_|k_adam_multi|k_sumsq_multi|k_bg_bwd' -s 14 -c 7 -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"; tail -3 gpurun_out/ncu2.log; ls -la gpurun_out/
