#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
python tools/render_order_prof.py > gpurun_out/rop_plain.log 2>&1; echo "plain rc=$?"; tail -4 gpurun_out/rop_plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_ngp_hash_fwd' --launch-skip 3 --launch-count 1 -f -o gpurun_out/r2_hash_fwd python tools/render_order_prof.py > gpurun_out/rop_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
