#!/bin/bash
# Round 2, GPU call N (4 GPUs): the bench line at N = 4 (the only rank count of the driver's scaling run not measured yet this round).
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_n4.log 2>&1; echo bench4=$?
tail -1 gpurun_out/bench_n4.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d.get(k) for k in ('value','ms_per_step','checked','n_gpus')})); print(json.dumps({k:(d['c5'] or {}).get(k) for k in ('ms_per_step','speedup_vs_n1','checked')})); print(json.dumps((d.get('gcn_epoch') or {}).get('ms')))"
