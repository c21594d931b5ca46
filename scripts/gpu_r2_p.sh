#!/bin/bash
# Round 2, GPU call P (1 GPU): sampling tests + benchmark after the register fast path; ncu --set full of the all-rows launch.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sampling_gpu.py tests/test_packed_registry.py -m gpu -x -q > gpurun_out/pytest_p.log 2>&1; echo pytest=$?; tail -2 gpurun_out/pytest_p.log
CONFIGS_OUT=configs_sampling.jsonl timeout 600 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_p.log 2>&1; echo sampling=$?
head -3 gpurun_out/sampling_p.log
CONFIGS_OUT=configs_sampling_p2.jsonl timeout 900 ncu --set full --clock-control none --import-source on -k regex:neighbor_sample_kernel -s 100 -c 1 \
    -o gpurun_out/prof_sampling -f python benchmarks/configs_bench.py sampling > gpurun_out/ncu_full_p.log 2>&1; echo ncu_full=$?
ncu -i gpurun_out/prof_sampling.ncu-rep --page raw --csv > gpurun_out/prof_sampling_raw.csv 2>/dev/null; echo raw=$?
