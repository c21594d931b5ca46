#!/bin/bash
# Memory and race check of the CUDA kernel sources on the HOST (no GPU): builds tests/native/cuda_emu with
# AddressSanitizer / ThreadSanitizer and runs the kernel parity bodies of tests/test_kernels_host_emu.py on it.
# Every CUDA thread is a host thread there, so an out-of-bounds access of any lane aborts under ASan and an
# unsynchronised access between lanes of a block (missing __syncthreads / __syncwarp) is a TSan report -
# the host counterpart of compute-sanitizer memcheck / racecheck.
#   scripts/emu_sanitize.sh address|thread
set -e
KIND=${1:-address}
SEL="spmm or csr or sddmm or gather or softmax or bucketing or sampling"
if [ "$KIND" = address ]; then
  export LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:halt_on_error=1
else
  export LD_PRELOAD=$(gcc -print-file-name=libtsan.so)
  export TSAN_OPTIONS="report_signal_unsafe=0 halt_on_error=0 exitcode=66 log_path=/tmp/dglb200_tsan"
  rm -f /tmp/dglb200_tsan.*
fi
DGLB200_EMU_SANITIZE=$KIND python -m pytest tests/test_kernels_host_emu.py -q -k "$SEL"
if [ "$KIND" = thread ]; then
  n=$(cat /tmp/dglb200_tsan.* 2>/dev/null | grep -c "WARNING: ThreadSanitizer" || true)
  echo "ThreadSanitizer reports: $n"
  [ "$n" = 0 ]
fi
