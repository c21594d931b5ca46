#!/bin/bash
# Round 2, GPU call X (8 GPUs): C2 all-gather with unequal slab widths (a narrow first slab lands sooner).
mkdir -p gpurun_out
for S in "" "64,192" "64,64,128" "32,96,128" "64,64,64,64"; do
  DGL_B200_AG_SLABS="$S" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29591 \
      bench.py --gpus 8 --steps 20 --warmup 5 --no-c5 --no-gcn --no-e2e --no-sampling > gpurun_out/bench_slabs.log 2>&1
  tail -1 gpurun_out/bench_slabs.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); c=d['config']; print('slabs=[$S]', round(d['ms_per_step'],3), 'op', round(c['op_ms_rank0'],3), 'exposed', round(c['exposed_exchange_ms_rank0'],3), 'checked', d['checked'])"
done
