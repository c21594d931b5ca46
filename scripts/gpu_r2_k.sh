#!/bin/bash
# Round 2, GPU call K (1 GPU): -m gpu suite; sampling benchmark after the bitmap / discovery rewrite, + its launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_k.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_k.log
CONFIGS_OUT=configs_sampling.jsonl timeout 900 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_k.log 2>&1; echo sampling=$?
tail -6 gpurun_out/sampling_k.log
CONFIGS_OUT=configs_sampling_ncu.jsonl timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"sample|scan|bitmap|place|csr_row" -c 400 --csv \
    --log-file gpurun_out/sampling_launches.csv python benchmarks/configs_bench.py sampling > gpurun_out/ncu_sampling_k.log 2>&1; echo ncu=$?
