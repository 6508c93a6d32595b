mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check.log 2>&1; echo "mgpu check rc=$?"; grep -v "^W\|^\*\*\*" gpurun_out/mgpu_check.log | tail -15
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench 2gpu rc=$?"; tail -3 gpurun_out/bench_2gpu.err
python -c "
import json
for l in open('gpurun_out/bench_2gpu.json'):
    if l.startswith('{'):
        d=json.loads(l); print('2gpu train', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'render', d['render'])
"
