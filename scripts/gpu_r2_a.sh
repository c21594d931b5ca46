#!/bin/bash
# Round 2, GPU call A (1 GPU): parity tests, smoke, the bench line (C2 headline + C5 record, self-checked),
# per-config timings, then ncu: C5 launch list + full capture of the C5 SpMM, launch lists of kernel
# families (3) and (4).  Every ncu pass runs directly after the same command exited 0 without ncu.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke=$?
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo bench=$?
tail -1 gpurun_out/bench.log | cut -c1-3000
timeout 900 python benchmarks/configs_bench.py > gpurun_out/configs.log 2>&1; echo configs=$?
C5="python bench.py --workload c5 --steps 2 --warmup 3 --no-e2e --no-cpu --no-gcn --no-check"
timeout 300 $C5 > gpurun_out/c5_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_" -c 40 --csv \
    --log-file gpurun_out/c5_launches.csv $C5 > gpurun_out/c5_ncu_list.log 2>&1; echo c5_list=$?
timeout 300 $C5 > gpurun_out/c5_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 4 -c 1 \
    -o gpurun_out/prof_c5_spmm -f $C5 > gpurun_out/c5_ncu_full.log 2>&1; echo c5_full=$?
K3="python benchmarks/configs_bench.py c3 c3big"
CONFIGS_OUT=k3.jsonl timeout 600 $K3 > gpurun_out/k3_plain.log 2>&1 && \
CONFIGS_OUT=k3_ncu.jsonl timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none \
    -k regex:"sddmm|softmax|gat_" -c 120 --csv --log-file gpurun_out/k3_launches.csv $K3 > gpurun_out/k3_ncu.log 2>&1; echo k3_list=$?
K4="python benchmarks/configs_bench.py c2"
CONFIGS_OUT=k4_ncu.jsonl timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"radix|csr_|scan|bucket|unique|hist" -c 150 --csv --log-file gpurun_out/k4_launches.csv $K4 > gpurun_out/k4_ncu.log 2>&1; echo k4_list=$?
ls -la gpurun_out | head -40
