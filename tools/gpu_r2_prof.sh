#!/bin/bash
# round-2 evidence: bench line, ncu launch list of the same eager step, per-kernel DRAM traffic / pipe metrics for the training
# step and for one 1080p render block, and full captures of the four heaviest kernels (ncu replays each launch ~40 times)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct,launch__registers_per_thread,launch__grid_size,launch__block_size"
python bench.py --steps 30 --warmup 5 > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err; echo "bench rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-secondary"
$CMD > gpurun_out/p_plain.log 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/p_launches.csv $CMD > gpurun_out/p_ncu1.log 2>&1; echo "launch list rc=$?"
ncu --metrics $M --clock-control none --launch-skip 500 -c 200 --csv --log-file gpurun_out/p_train_kernels.csv $CMD > gpurun_out/p_ncu2.log 2>&1; echo "train metrics rc=$?"
python tools/render_order_prof.py > gpurun_out/p_plain2.log 2>&1; echo "render plain rc=$?"
ncu --metrics $M --clock-control none --csv --log-file gpurun_out/p_render_kernels.csv python tools/render_order_prof.py > gpurun_out/p_ncu3.log 2>&1; echo "render metrics rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_mlp_fwd_tc' --launch-skip 3 --launch-count 1 -f -o gpurun_out/r2_fused_fwd python tools/render_order_prof.py > gpurun_out/p_ncu4.log 2>&1; echo "fused full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_hash_bwd|k_ngp_mlp_bwd_tc|k_adam_multi|k_ngp_hash_fwd' --launch-skip 20 --launch-count 4 -f -o gpurun_out/r2_train_top $CMD > gpurun_out/p_ncu5.log 2>&1; echo "train full rc=$?"
ls -la gpurun_out | tail -14
