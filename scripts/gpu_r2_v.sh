#!/bin/bash
# Round 2, GPU call V (8 GPUs): the bench line at N = 8 on the final code.
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 \
    bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8.log 2>&1; echo bench8=$?
tail -1 gpurun_out/bench_n8.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d.get(k) for k in ('value','ms_per_step','checked','n_gpus')})); print(json.dumps({k:(d['c5'] or {}).get(k) for k in ('ms_per_step','speedup_vs_n1','checked')})); print(json.dumps({k:(d['c5_bf16_wire'] or {}).get(k) for k in ('ms_per_step','speedup_vs_n1','checked')})); print(json.dumps((d.get('gcn_epoch') or {}).get('ms')), json.dumps(d['e2e']))"
