"""TEST INFRASTRUCTURE ONLY - builds libdgl_b200_emu.so: the UNMODIFIED CUDA sources of
dgl-1_b200/csrc compiled for the host against the execution model of cuda_runtime.h in this
directory, exporting the same C ABI as libdgl_b200.so (include/dgl_b200.h).

The only source transformation is the launch syntax, which is not C++:
    kernel<targs><<<grid, block, smem, stream>>>(args)
 -> ::dglb200_emu::launch(grid, block, [&]() { kernel<targs>(args); })
packed.cu / packed_graph.cu (PackedFunc registration against a loaded libdgl.so) are part of the build:
their entry points drive the emulated kernels when a test registers them into the compiled reference.
"""
import hashlib
import os
import re
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(HERE)))
CSRC = os.path.join(ROOT, "dgl-1_b200", "csrc")
SKIP = set()


def _match(text, i, open_c, close_c):
    """Index just past the bracket that closes the one at text[i]."""
    depth = 0
    while i < len(text):
        c = text[i]
        if c == open_c:
            depth += 1
        elif c == close_c:
            depth -= 1
            if depth == 0:
                return i + 1
        i += 1
    raise ValueError("unbalanced %s%s" % (open_c, close_c))


def _split_top(s):
    """Split a comma list at nesting depth 0."""
    out, depth, cur = [], 0, ""
    for c in s:
        if c in "([{<":
            depth += 1
        elif c in ")]}>":
            depth -= 1
        if c == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += c
    out.append(cur)
    return out


def rewrite_launches(text):
    out, pos = "", 0
    while True:
        k = text.find("<<<", pos)
        if k < 0:
            return out + text[pos:]
        line_start = text.rfind("\n", 0, k) + 1
        if "//" in text[line_start:k]:                      # a comment that mentions the syntax
            out += text[pos:k + 3]
            pos = k + 3
            continue
        # kernel expression: identifier, optionally followed by balanced template arguments
        j = k
        if text[j - 1] == ">":
            depth, j = 0, k - 1
            while True:
                if text[j] == ">":
                    depth += 1
                elif text[j] == "<":
                    depth -= 1
                    if depth == 0:
                        break
                j -= 1
        m = re.search(r"[A-Za-z_][A-Za-z_0-9:]*$", text[:j])
        start = m.start()
        kernel = text[start:k]
        e = text.index(">>>", k)
        cfg = _split_top(text[k + 3:e].replace("\\\n", " "))
        a0 = e + 3
        while text[a0] in " \\\n":
            a0 += 1
        assert text[a0] == "(", text[k - 40:e + 40]
        a1 = _match(text, a0, "(", ")")
        args = text[a0:a1]
        out += text[pos:start]
        out += "::dglb200_emu::launch(%s, %s, [&]() { %s%s; })" % (cfg[0].strip(), cfg[1].strip(), kernel, args)
        pos = a1


def build(out_dir=None, sanitize=None):
    """Returns the path of the host-emulation library, (re)building it when the sources changed.
    sanitize="address": AddressSanitizer build (every out-of-bounds access of a kernel aborts; the
    process that loads it must preload libasan, see tests/test_kernels_host_emu.py)."""
    sanitize = sanitize or os.environ.get("DGLB200_EMU_SANITIZE") or None
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh")))
    h = hashlib.sha1()
    h.update(str(sanitize).encode())
    for f in srcs + ["cuda_runtime.h", "cuda_bf16.h", "build.py"]:
        p = os.path.join(CSRC, f) if f in srcs else os.path.join(HERE, f)
        h.update(open(p, "rb").read())
    out_dir = out_dir or os.path.join(tempfile.gettempdir(), "dgl_b200_emu_" + h.hexdigest()[:12])
    lib = os.path.join(out_dir, "libdgl_b200_emu.so")
    if os.path.exists(lib):
        return lib
    if shutil.which("g++") is None:
        raise RuntimeError("no g++")
    work = out_dir + ".tmp%d" % os.getpid()
    os.makedirs(work, exist_ok=True)
    cpps = []
    for f in srcs:
        text = rewrite_launches(open(os.path.join(CSRC, f)).read())
        name = f[:-3] + ".cpp" if f.endswith(".cu") else f
        with open(os.path.join(work, name), "w") as fh:
            fh.write(text)
        if f.endswith(".cu") and f not in SKIP:
            cpps.append(os.path.join(work, name))
    flags = ["g++", "-std=c++20", "-O1", "-fPIC", "-pthread", "-w", "-DDGLB200_HOST_EMU",
             "-I", HERE, "-I", work, "-I", os.path.join(ROOT, "include")]
    san = ["-fsanitize=%s" % sanitize, "-fno-omit-frame-pointer", "-g"] if sanitize else []
    flags += san
    tls = os.path.join(work, "emu_state.cpp")
    with open(tls, "w") as fh:
        fh.write('#include "cuda_runtime.h"\nthread_local dim3 threadIdx;\ndim3 blockIdx, blockDim, gridDim;\n')
    objs = []
    procs = []
    for c in cpps + [tls]:
        o = c[:-4] + ".o"
        objs.append(o)
        procs.append((c, subprocess.Popen(flags + ["-c", c, "-o", o],
                                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for c, p in procs:
        log = p.communicate()[0].decode()
        if p.returncode != 0:
            raise RuntimeError("host build of %s failed:\n%s" % (os.path.basename(c), log[-6000:]))
    # -Bsymbolic: the real libdgl_b200.so may be loaded RTLD_GLOBAL in the same process and defines the
    # same symbols; this library must bind its internal references to its own definitions
    subprocess.check_call(["g++", "-shared", "-pthread", "-Wl,-Bsymbolic"] + san + ["-o",
                           os.path.join(work, "libdgl_b200_emu.so")] + objs + ["-ldl"])
    if os.path.isdir(out_dir):
        shutil.rmtree(work, ignore_errors=True)
    else:
        os.rename(work, out_dir)
    return lib


if __name__ == "__main__":
    print(build())
