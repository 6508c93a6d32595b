#!/bin/bash
# round 2: multi-GPU parity (tools/mgpu_check.py) + the bench line at N = number of visible GPUs
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
timeout 300 python -m pytest tests/test_gpu_step.py tests/test_gpu_multi.py -q -m gpu > gpurun_out/m_tests.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/m_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py > gpurun_out/mgpu_check_${N}gpu.log 2>&1; echo "mgpu check rc=$?"; grep -v "^W\|^\*\*\*" gpurun_out/mgpu_check_${N}gpu.log | tail -15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"; grep -v "^W\|^\*\*\*" gpurun_out/bench_${N}gpu.err | tail -25
python - <<PY
import json
for l in open('gpurun_out/bench_${N}gpu.json'):
    if l.startswith('{'):
        d=json.loads(l)
        print('train', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
        print('parity', json.dumps(d.get('parity_check'))[:1500])
        for k in ('render_c5','train_c2','render_800','adapt_c4'):
            v=d.get(k)
            if isinstance(v,dict): v={a:b for a,b in v.items() if a not in('kernels','config','roofline','e2e')}
            print(k, json.dumps(v)[:400])
        for k,v in d['kernels'].items():
            print(f"{k:18s} {v['ms_per_step']*1000:9.1f} us  calls {v['calls_per_step']:.1f}  frac {v.get('frac')}")
PY
