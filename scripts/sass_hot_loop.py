"""Static evidence for DESIGN.md section 3: instruction mix of the headline SpMM kernel
(spmm_kernel<float, VEC=4, G=16, TILES=1, copy_src, sum, NT=128>) from the built library's SASS.

    python scripts/sass_hot_loop.py > profiles/r01_sass_spmm_hot_loop.txt

Needs only cuobjdump (no GPU).  Prints the whole-kernel opcode histogram, every backward branch
(loop) with its length, and the listing of the predicate-free FULL-block loop."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dgl-1_b200", "dgl_b200", "lib", "libdgl_b200.so")
KERNEL = "_ZN7dglb20011spmm_kernelIfLi4ELi16ELi1ELi0ELi0ELi0ELi128ELi0EEEvNS_8SpmmArgsE"


def main():
    sass = subprocess.check_output(["cuobjdump", "-sass", LIB]).decode()
    blk = next(b for b in sass.split("\t\tFunction : ") if b.startswith(KERNEL))
    ops = []
    for ln in blk.splitlines():
        m = re.search(r"/\*([0-9a-f]{4})\*/\s+((?:@!?U?P\d+\s+)?[A-Z0-9_.]+.*?);", ln)
        if m:
            ops.append((int(m.group(1), 16), m.group(2).strip()))
    name = lambda o: re.sub(r"^@!?U?P\d+\s+", "", o).split()[0]  # noqa: E731
    print("kernel: spmm_kernel<float,4,16,1,copy_src,sum,0,128,any>  (%s)" % KERNEL)
    print("instructions: %d" % len(ops))
    print("opcode histogram:", collections.Counter(name(o).split(".")[0] for _, o in ops).most_common())
    print("memory ops:", collections.Counter(name(o) for _, o in ops if name(o).startswith(("LDG", "STG", "LDS", "STS"))).most_common())
    loops = []
    for a, o in ops:
        if name(o).startswith("BRA"):
            m = re.search(r"0x([0-9a-f]+)", o)
            if m and int(m.group(1), 16) < a:
                loops.append((int(m.group(1), 16), a))
    print("loops (backward branches):", [(hex(lo), hex(hi), (hi - lo) // 16 + 1) for lo, hi in loops])
    # the FULL-block loop is the shortest loop that holds LDG.E.128 and no predicated loads
    cand = []
    for lo, hi in loops:
        body = [o for a, o in ops if lo <= a <= hi]
        if sum(1 for o in body if "LDG.E.128" in o) >= 8:
            cand.append((len(body), lo, hi))
    if not cand:
        return 0
    _, lo, hi = min(cand)
    body = [(a, o) for a, o in ops if lo <= a <= hi]
    mix = collections.Counter(name(o).split(".")[0] for _, o in body)
    n_ldg = sum(1 for _, o in body if "LDG.E.128" in o)
    print("\nFULL-block loop %s..%s: %d instructions, %d LDG.E.128 (= edge-lanes per trip on the longest path)"
          % (hex(lo), hex(hi), len(body), n_ldg))
    print("mix:", mix.most_common())
    for a, o in body:
        print("  %04x  %s" % (a, o))
    return 0


if __name__ == "__main__":
    sys.exit(main())
