#!/bin/bash
# Round 2, GPU call F (2 GPUs): the full -m gpu suite (2-rank transports, fused GCN, no-alloc entry point),
# the N=1 bench line (fused GCN epoch, C5 record) and the N=2 line (C5 on both wires).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_f.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_f.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_f_n1.log 2>gpurun_out/bench_f_n1.err; echo bench1=$?
tail -3 gpurun_out/bench_f_n1.err | cut -c1-300
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29501 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_f_n2.log 2>gpurun_out/bench_f_n2.err; echo bench2=$?
tail -3 gpurun_out/bench_f_n2.err | cut -c1-300
