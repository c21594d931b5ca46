#!/bin/bash
# Round 2, GPU call H (1 GPU): full -m gpu suite, per-config timings (after the kernel 3/4 rewrites),
# compute-sanitizer memcheck + racecheck over a kernel-parity subset, ncu launch list + full capture of the
# C2 headline kernel (each ncu pass directly after the same command exited 0 without ncu).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_h.log 2>&1; echo pytest=$?; tail -2 gpurun_out/pytest_h.log
timeout 900 python benchmarks/configs_bench.py > gpurun_out/configs_h.log 2>&1; echo configs=$?
SUB="tests/test_gpu_spmm.py tests/test_gpu_sddmm_bucketing.py tests/test_gpu_csr_build.py tests/test_fused_gcn_gpu.py tests/test_zz_no_alloc.py"
KSEL="known_answers or (copy_src_sum and 64-) or (src_mul_edge and 64-) or max_with_argmax or bf16 or backward_x or accumulate or edge_softmax or gat_aggregate or degree_bucketing or unique or csr_mutable or fused or static_scratch"
timeout 1500 compute-sanitizer --tool memcheck --target-processes all --error-exitcode 86 python -m pytest $SUB -m gpu -q -x -k "$KSEL" > gpurun_out/sanitizer_memcheck.log 2>&1; echo memcheck=$?
tail -4 gpurun_out/sanitizer_memcheck.log
timeout 1500 compute-sanitizer --tool racecheck --target-processes all --error-exitcode 86 python -m pytest $SUB -m gpu -q -x -k "$KSEL" > gpurun_out/sanitizer_racecheck.log 2>&1; echo racecheck=$?
tail -4 gpurun_out/sanitizer_racecheck.log
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-gcn --no-c5 --no-check"
$BENCH > gpurun_out/plain_h.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"spmm_|plan_|csr_|radix_|scan_" -c 200 --csv \
    --log-file gpurun_out/c2_launches.csv $BENCH > gpurun_out/ncu_list_h.log 2>&1; echo ncu_list=$?
$BENCH > gpurun_out/plain_h2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 6 -c 1 \
    -o gpurun_out/prof_c2_spmm -f $BENCH > gpurun_out/ncu_full_h.log 2>&1; echo ncu_full=$?
