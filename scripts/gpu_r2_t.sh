#!/bin/bash
# Round 2, GPU call T (1 GPU): the round-end sequence on the final code - -m gpu suite, smoke(), bench.py default line, --impl reference (short).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo pytest=$?; tail -2 gpurun_out/pytest_t.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/smoke_t.log 2>&1; echo smoke=$?; tail -1 gpurun_out/smoke_t.log
timeout 900 python bench.py > gpurun_out/bench_t.log 2>&1; echo bench=$?
tail -1 gpurun_out/bench_t.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({k:d[k] for k in ('value','ms_per_step','checked','gpu_launches')})); print(json.dumps(d['roofline'])[:300]); print(json.dumps(d['neighbor_sampling'])[:400]); print(json.dumps(d['e2e']))"
