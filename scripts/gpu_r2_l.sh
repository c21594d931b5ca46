#!/bin/bash
# Round 2, GPU call L (1 GPU): sampling tests + benchmark after moving the whole call into dglb200_sample_subgraph.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sampling_gpu.py tests/test_abi_symbols.py -m gpu -x -q > gpurun_out/pytest_l.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_l.log
CONFIGS_OUT=configs_sampling.jsonl timeout 900 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_l.log 2>&1; echo sampling=$?
tail -6 gpurun_out/sampling_l.log
