#!/bin/bash
# Round 2, GPU call B (N GPUs, default 2): the peer transports on real NVLink - the 2-rank pytest, the
# bench line at N ranks (C2 headline + C5 record, both self-checked), the C5 knob sweep.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_round2.py -x -q -m gpu -k two_ranks > gpurun_out/pytest_2rank.log 2>&1; echo pytest2=$?; tail -5 gpurun_out/pytest_2rank.log
timeout 900 $TR --master-port 29501 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>gpurun_out/bench_n$N.err; echo bench=$?
tail -1 gpurun_out/bench_n$N.log | cut -c1-4000
tail -5 gpurun_out/bench_n$N.err
timeout 1200 $TR --master-port 29502 scripts/c5_sweep.py ${SWEEP_ARGS:-} > gpurun_out/sweep_n$N.log 2>gpurun_out/sweep_n$N.err; echo sweep=$?
grep '^{' gpurun_out/sweep_n$N.log | cut -c1-400
tail -5 gpurun_out/sweep_n$N.err
