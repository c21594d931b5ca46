#!/bin/bash
# Round 2, GPU call Q (1 GPU): kernel family (3) at C2 size with edge data in slot order; the autograd GAT chain both ways;
# ncu raw of the final sampling kernel; edge-ops tests.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_uopv_edge_ops_gpu.py tests/test_gpu_sddmm_bucketing.py -m gpu -x -q > gpurun_out/pytest_q.log 2>&1; echo pytest=$?; tail -2 gpurun_out/pytest_q.log
CONFIGS_OUT=configs_c3big.jsonl timeout 900 python benchmarks/configs_bench.py c3big > gpurun_out/c3big_q.log 2>&1; echo c3big=$?
grep -o '"op": "[^"]*", "us": [0-9.]*' gpurun_out/c3big_q.log
CONFIGS_OUT=configs_gatchain.jsonl timeout 900 python benchmarks/configs_bench.py gatchain > gpurun_out/gatchain_q.log 2>&1; echo gatchain=$?
grep -o '"op": "[^"]*", "us": [0-9.]*' gpurun_out/gatchain_q.log; tail -3 gpurun_out/gatchain_q.log | grep -i error
CONFIGS_OUT=configs_sampling_q.jsonl timeout 600 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_q.log 2>&1 && \
CONFIGS_OUT=configs_sampling_q2.jsonl timeout 900 ncu --set full --clock-control none --import-source on -k regex:neighbor_sample_kernel -s 100 -c 1 \
    -o gpurun_out/prof_sampling -f python benchmarks/configs_bench.py sampling > gpurun_out/ncu_full_q.log 2>&1; echo ncu_full=$?
ncu -i gpurun_out/prof_sampling.ncu-rep --page raw --csv > gpurun_out/prof_sampling_raw.csv 2>/dev/null; echo raw=$?
