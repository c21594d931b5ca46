#!/bin/bash
# Round 2, GPU call J (1 GPU): CUDA-graph capture of the C1 / C3 / C4 training steps (one process per config).
mkdir -p gpurun_out
for c in C1 C3 C4; do
  CAPTURE_ONLY=$c CONFIGS_OUT=configs_capture_$c.jsonl timeout 600 python benchmarks/configs_bench.py capture > gpurun_out/capture_$c.log 2>&1; echo capture_$c=$?
  grep -v Warning gpurun_out/capture_$c.log | tail -12
done
