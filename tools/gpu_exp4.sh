mkdir -p gpurun_out
M="gpu__time_duration.sum,lts__t_sectors_op_red.sum,lts__t_requests_op_red.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_requests_pipe_lsu_mem_global_op_red.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum,l1tex__data_pipe_lsu_wavefronts_mem_global_op_red.sum,smsp__inst_executed.sum,smsp__inst_executed_op_global_red.sum,lts__t_sectors_srcunit_tex_op_red.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sectors_op_red_lookup_miss.sum,lts__t_sectors_op_red_lookup_hit.sum"
ncu --metrics $M --clock-control none -k regex:'^k$' -s 36 -c 1 --csv --log-file gpurun_out/ncu_replica.csv ./scratch/red_bench2 > gpurun_out/ncu_replica.log 2>&1
echo "replica rc=$?"
ACN_HASH_BWD=0 ncu --metrics $M --clock-control none -k regex:'k_ngp_hash_bwd' -s 3 -c 1 --csv --log-file gpurun_out/ncu_real.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-graph > gpurun_out/ncu_real.log 2>&1
echo "real rc=$?"
