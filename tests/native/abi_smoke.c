/* A plain-C client of the drop-in boundary (include/dgl_b200.h): what a cgo / JNI / ctypes binding
 * sees.  Built and run by tests/test_abi_host.py with gcc -std=c99 -pedantic; needs no GPU: it only
 * exercises the calls that validate their arguments before touching CUDA. */
#include <stdio.h>
#include <string.h>

#include "dgl_b200.h"

int main(void) {
  size_t bytes = 0;
  if (dglb200_abi_version() != 1) return 1;
  if (dglb200_csr_build_workspace_bytes(10, 100, &bytes) != 0 || bytes == 0) return 2;
  if (dglb200_spmm_plan_workspace_bytes(10, &bytes) != 0) return 3;
  /* errors: -1 + a thread-local message, never abort (cf. DGLGetLastError) */
  if (dglb200_set_tuning("no_such_knob", 1) != -1) return 4;
  if (strstr(dglb200_last_error(), "no_such_knob") == NULL) return 5;
  if (dglb200_spmm_csr(-1, 0, 0, NULL, NULL, NULL, 0, 0, 0, NULL, 4, NULL, 1, NULL, NULL, NULL, NULL) != -1)
    return 6;
  if (strstr(dglb200_last_error(), "bad sizes") == NULL) return 7;
  if (dglb200_sddmm_csr(1, 1, NULL, NULL, NULL, 0, 0, NULL, NULL, 1, 1, NULL, NULL) != -1) return 8;
  if (dglb200_bucket_capacity(100) <= 0) return 9;
  printf("abi ok, launches so far: %lld\n", (long long)dglb200_launch_count());
  return 0;
}
