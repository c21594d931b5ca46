#!/bin/bash
# Round 2, GPU call M (1 GPU): full -m gpu suite (sampler overrides in libdgl, CUDA-graph capture test), sampling
# benchmark with the sync-free bounded hops, its launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_m.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_m.log
CONFIGS_OUT=configs_sampling.jsonl timeout 900 python benchmarks/configs_bench.py sampling > gpurun_out/sampling_m.log 2>&1; echo sampling=$?
tail -6 gpurun_out/sampling_m.log
CONFIGS_OUT=configs_sampling_ncu.jsonl timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"sample|scan|bitmap|place" -c 400 --csv \
    --log-file gpurun_out/sampling_launches.csv python benchmarks/configs_bench.py sampling > gpurun_out/ncu_sampling_m.log 2>&1; echo ncu=$?
timeout 900 python bench.py > gpurun_out/bench_m.log 2>&1; echo bench=$?; tail -1 gpurun_out/bench_m.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d[k] for k in ('value','ms_per_step','checked','neighbor_sampling')}))"
