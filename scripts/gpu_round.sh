#!/bin/bash
# One GPU-box round: parity tests, smoke, bench, then the ncu launch list and one full capture of
# the dominant kernel (each ncu pass directly after the same command exited 0 without ncu).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke=$?
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>&1; echo bench=$?
tail -1 gpurun_out/bench.log
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$BENCH > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none \
    -k regex:"spmm_|plan_|csr_|radix_|scan_|to_u32|gather_u32" -c 200 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_list.log 2>&1; echo ncu_list=$?
$BENCH > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 6 -c 2 \
    -o gpurun_out/prof_spmm -f $BENCH > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
