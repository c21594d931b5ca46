#!/bin/bash
# Round 2, GPU call W (1 GPU): launch list of kernel family (3) at C2 size on the final code (after the plain run).
mkdir -p gpurun_out
CONFIGS_OUT=configs_c3big_w.jsonl timeout 900 python benchmarks/configs_bench.py c3big > gpurun_out/c3big_w.log 2>&1 && \
CONFIGS_OUT=configs_c3big_w2.jsonl timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"sddmm|softmax|gat_aggregate" -c 200 --csv --log-file gpurun_out/k3_c2size_launches.csv python benchmarks/configs_bench.py c3big > gpurun_out/ncu_w.log 2>&1; echo ncu=$?
