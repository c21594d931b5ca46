#!/bin/bash
# Round 2, GPU call I (1 GPU): full -m gpu suite (adds neighbour sampling), the sampling benchmark against the
# compiled reference on the host cores, a launch list of that benchmark, bench.py's default line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_i.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_i.log
CONFIGS_OUT=configs_sampling.jsonl timeout 900 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_i.log 2>&1; echo sampling=$?
tail -6 gpurun_out/sampling_i.log
timeout 900 python bench.py > gpurun_out/bench_i.log 2>&1; echo bench=$?; tail -1 gpurun_out/bench_i.log
