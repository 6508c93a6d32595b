#!/bin/bash
# round-1 v6 evidence: bench line, ncu launch list of the same command, one full capture of the routing / exchange / top kernels
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_all.log 2>&1; echo "gpu tests rc=$?"; tail -2 gpurun_out/gpu_all.log
python bench.py --steps 20 --warmup 3 > gpurun_out/b1.json 2> gpurun_out/b1.err; echo "bench rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-c3 --no-graph"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
python tools/route_bench.py 1 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_route_' --launch-skip 4 --launch-count 4 -f -o gpurun_out/prof6_route python tools/route_bench.py 1 > gpurun_out/ncu2.log 2>&1
echo "route full rc=$?"
python -m pytest tests/test_gpu_parity.py -q -k peer_exchange > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:'k_a2a_|k_p2p_' --csv --log-file gpurun_out/a2a_kernels.csv python -m pytest tests/test_gpu_parity.py -q -k peer_exchange > gpurun_out/ncu3.log 2>&1
echo "a2a rc=$?"
ls -la gpurun_out | tail -12
