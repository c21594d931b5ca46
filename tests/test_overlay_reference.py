"""dgl_b200.overlay executed, not described: the reference's own Python (imported from /root/reference,
build container only) runs its builtin message-passing path through the engine's C ABI after
overlay.install() patched the four call sites SURVEY.md 8(b) names.  The driver runs in a subprocess (the
patches and the PackedFunc overrides are process-wide) and replays the reference's
tests/compute/test_specialization.py + test_function.py + test_sampler.py under the overlay."""
import os
import shutil
import subprocess
import sys

import pytest

from oracle import ref_ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.isdir("/root/reference/python/dgl") and ref_ffi.available()


def _run(mode):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "overlay_driver.py"), mode],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=1200)
    out = p.stdout.decode("utf-8", "replace")
    ran = 14
    assert p.returncode == 0 and "OVERLAY_OK" in out and "REFERENCE_TESTS_OK %d" % ran in out, out[-4000:]
    assert "OVERLAY_CASES_OK 48 cases" in out and "REFERENCE_SAMPLER_TESTS_OK 5" in out


@pytest.mark.skipif(not HAVE_REF, reason="needs /root/reference/python and oracle/_ref/libdgl.so (build container)")
@pytest.mark.timeout(1500)
def test_reference_update_all_through_overlay_on_emulated_kernels():
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    _run("emu")


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE_REF, reason="the reference's Python does not travel to the GPU box")
def test_reference_update_all_through_overlay_on_the_gpu():
    _run("gpu")
