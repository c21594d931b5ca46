#!/bin/bash
# Round 2, GPU call C (2 GPUs): full parity suite incl. the 2-rank transports, kernel (3)/(4) timings after
# the rewrites, a short N=2 bench (own-block-first all-gather, mixed-precision halo).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_c.log 2>&1; echo pytest=$?; tail -4 gpurun_out/pytest_c.log
CONFIGS_OUT=configs_c.jsonl timeout 900 python benchmarks/configs_bench.py c3 c3big c2 > gpurun_out/configs_c.log 2>&1; echo configs=$?
grep -E "softmax|csr_build|bucketing|unique|sddmm|GAT" gpurun_out/configs_c.log | cut -c1-200
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29501 bench.py --gpus 2 --steps 10 --warmup 3 --no-gcn > gpurun_out/bench_c_n2.log 2>gpurun_out/bench_c_n2.err; echo bench=$?
tail -1 gpurun_out/bench_c_n2.log | cut -c1-1500; tail -3 gpurun_out/bench_c_n2.err
DGL_B200_WIRE=bf16 timeout 900 $TR --master-port 29503 bench.py --gpus 2 --workload c5 --steps 6 --warmup 3 --no-gcn --no-e2e > gpurun_out/bench_c_n2_bf16.log 2>gpurun_out/bench_c_n2_bf16.err; echo bench_bf16=$?
tail -1 gpurun_out/bench_c_n2_bf16.log | cut -c1-1200; tail -3 gpurun_out/bench_c_n2_bf16.err
