#!/bin/bash
# Round 2, GPU call R (1 GPU): SDDMM DOT in L2-resident head slabs at C2 size; the autograd GAT chain again; sddmm tests.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_uopv_edge_ops_gpu.py tests/test_gpu_sddmm_bucketing.py -m gpu -x -q > gpurun_out/pytest_r.log 2>&1; echo pytest=$?; tail -2 gpurun_out/pytest_r.log
CONFIGS_OUT=configs_c3big.jsonl timeout 900 python benchmarks/configs_bench.py c3big > gpurun_out/c3big_r.log 2>&1; echo c3big=$?
grep -o '"op": "[^"]*", "us": [0-9.]*' gpurun_out/c3big_r.log
CONFIGS_OUT=configs_gatchain.jsonl timeout 900 python benchmarks/configs_bench.py gatchain > gpurun_out/gatchain_r.log 2>&1; echo gatchain=$?
grep -o '"op": "[^"]*", "us": [0-9.]*' gpurun_out/gatchain_r.log
