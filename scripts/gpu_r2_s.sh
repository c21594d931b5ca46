#!/bin/bash
# Round 2, GPU call S (1 GPU): ncu --set full of the SDDMM DOT launch at C2 size (8 heads x 32), after the plain run.
mkdir -p gpurun_out
CONFIGS_OUT=configs_c3big_s.jsonl timeout 900 python benchmarks/configs_bench.py c3big > gpurun_out/c3big_s.log 2>&1 && \
CONFIGS_OUT=configs_c3big_s2.jsonl timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"sddmm_dot_warp_kernel|sddmm_kernel" -s 17 -c 1 \
    -o gpurun_out/prof_sddmm_dot -f python benchmarks/configs_bench.py c3big > gpurun_out/ncu_full_s.log 2>&1; echo ncu_full=$?
ncu -i gpurun_out/prof_sddmm_dot.ncu-rep --page raw --csv > gpurun_out/prof_sddmm_dot_raw.csv 2>/dev/null; echo raw=$?
