#!/bin/bash
# Round 2, GPU call O (1 GPU): ncu --set full of the sampling kernel (the all-rows launch at fanout 10) after the same command exited 0 without ncu.
mkdir -p gpurun_out
CONFIGS_OUT=configs_sampling_o.jsonl timeout 600 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_o.log 2>&1 && \
CONFIGS_OUT=configs_sampling_o2.jsonl timeout 900 ncu --set full --clock-control none --import-source on -k regex:neighbor_sample_kernel -s 100 -c 1 \
    -o gpurun_out/prof_sampling -f python benchmarks/configs_bench.py sampling > gpurun_out/ncu_full_o.log 2>&1; echo ncu_full=$?
ncu -i gpurun_out/prof_sampling.ncu-rep --page raw --csv > gpurun_out/prof_sampling_raw.csv 2>/dev/null; echo raw=$?
