#!/bin/bash
# ncu metrics for the kernels the bench step does not launch: occupancy, episode routing, Voronoi, PE-MLP (config 0), peer-exchange
# (slabs local on one GPU), proposal resampling, barriers
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct,launch__registers_per_thread,launch__grid_size,launch__block_size"
T="tests/test_gpu_occ.py tests/test_gpu_episode.py::test_episode_router_one_million_rays_vs_oracle tests/test_gpu_voronoi.py::test_voronoi_masks_boxes_and_counts_at_scale_vs_oracle tests/test_gpu_pe.py tests/test_gpu_parity.py::test_peer_exchange_path_matches_the_fused_container_on_one_gpu tests/test_gpu_tc.py::test_proposal_resampling_vs_oracle_and_hierarchical_render tests/test_gpu_step.py::test_split_barrier_arrive_then_wait_reduces_and_replays_in_a_graph tests/test_gpu_step.py::test_p2p_plan_kernel_matches_peer_layout"
python -m pytest -q -m gpu $T > gpurun_out/x_plain.log 2>&1; echo "plain rc=$?"; tail -2 gpurun_out/x_plain.log
timeout 900 ncu --metrics $M --clock-control none -k regex:'^k_|acn::k_' --csv --log-file gpurun_out/x_aux_kernels.csv python -m pytest -q -m gpu $T > gpurun_out/x_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/x_ncu.log
wc -l gpurun_out/x_aux_kernels.csv
