#!/bin/bash
# Round 2, GPU call G (8 GPUs): the N=8 bench line (C2 headline, C5 on both wires, all self-checked) and a C5
# sweep over the push variants (thread stores / TMA stores / TMA both ways), chunk counts and wires.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
DGL_B200_TIMING=1 timeout 900 $TR --master-port 29501 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_g_n8.log 2>gpurun_out/bench_g_n8.err; echo bench8=$?
tail -2 gpurun_out/bench_g_n8.err | cut -c1-300
export C5_SWEEP_GRID='[[1,null,"signal","f32",0],[4,null,"signal-lsu","f32",0],[4,null,"signal-tma1","f32",0],[4,null,"signal","f32",0],[4,null,"signal","f32",64],[4,null,"signal","f32",296],[6,null,"signal","f32",0],[4,null,"push","f32",0],[4,null,"signal-lsu","bf16",0],[4,null,"signal-tma1","bf16",148],[4,null,"signal-tma1","bf16",0],[2,null,"signal-tma1","bf16",0],[6,null,"signal-tma1","bf16",0]]'
timeout 900 $TR --master-port 29502 scripts/c5_sweep.py > gpurun_out/sweep_g_n8.log 2>gpurun_out/sweep_g_n8.err; echo sweep=$?
tail -2 gpurun_out/sweep_g_n8.err | cut -c1-300
