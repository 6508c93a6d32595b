#!/bin/bash
# end of round 2: ncu metrics of the routing pack kernels after the decision-word change (16.7 M samples x 8 experts), then the
# acceptance run (smoke, whole GPU suite, bench line, reference arm)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct,launch__registers_per_thread,launch__grid_size,launch__block_size"
timeout 120 python tools/route_bench.py 3 > gpurun_out/l_route_plain.log 2>&1; echo "route plain rc=$?"; tail -1 gpurun_out/l_route_plain.log
timeout 300 ncu --metrics $M --clock-control none -k regex:'k_route' --csv --log-file gpurun_out/l_route_kernels.csv python tools/route_bench.py 2 > gpurun_out/l_ncu.log 2>&1; echo "route ncu rc=$?"
bash tools/gpu_r2_final.sh
