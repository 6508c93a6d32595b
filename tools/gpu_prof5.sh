mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_|k_adam_multi|k_sumsq_multi' -s 28 -c 6 -o gpurun_out/prof5 $CMD > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"; tail -3 gpurun_out/ncu2.log
M="gpu__time_duration.sum,smsp__inst_executed_op_global_red.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_red_lookup_hit.sum,lts__t_sectors_op_red_lookup_miss.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum"
ncu --metrics $M --clock-control none -k regex:'k_ngp_hash_bwd' -s 3 -c 1 --csv --log-file gpurun_out/hash_bwd_red.csv $CMD > gpurun_out/ncu3.log 2>&1
echo "red rc=$?"
ls -la gpurun_out/ | head -30
