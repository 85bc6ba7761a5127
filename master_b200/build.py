"""Builds master_b200/libb2pt.so (the C-ABI library of include/b2pt.h) with nvcc for sm_100a.

In-tree and explicit so the built .so travels to the GPU box with the repo snapshot.  No torch
types cross the ABI, so this is plain `nvcc -shared`, not a torch extension.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb2pt.so")
OBJ = os.path.join(HERE, "build")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # x86-64 semantics for the shading arithmetic: no silent FMA contraction, FTZ like main.cpp:70-71,
    # IEEE division and square root (never -use_fast_math); see csrc/common.cuh
    "--fmad=false", "-ftz=true", "-prec-div=true", "-prec-sqrt=true",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2",
    "--expt-relaxed-constexpr", "-diag-suppress", "177,550",
]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stamp(paths):
    h = hashlib.sha1()
    for p in sorted(paths):
        h.update(p.encode())
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_library(force=False, verbose=False, variant=None, extra_flags=()):
    """Compile every .cu under csrc/ and link libb2pt.so. Returns the library path.

    `variant` + `extra_flags` build an A/B copy (libb2pt_<variant>.so, own object directory) that tools select with
    B2PT_LIB=<path>; the product library is always the flag set above."""
    global OBJ, LIB
    flags = NVCC_FLAGS + list(extra_flags)
    if variant:
        OBJ = os.path.join(HERE, "build", variant)
        LIB = os.path.join(HERE, "libb2pt_%s.so" % variant)
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".hpp"))]
    headers.append(os.path.join(HERE, "..", "include", "b2pt.h"))
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs, procs = [], []
    for src in sources():
        path = os.path.join(CSRC, src)
        obj = os.path.join(OBJ, src[:-3] + ".o")
        stamp_file = obj + ".stamp"
        stamp = _stamp([path] + headers) + " ".join(extra_flags)
        objs.append(obj)
        if not force and os.path.exists(obj) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
            continue
        cmd = [nvcc] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
        procs.append((src, stamp_file, stamp, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    failed = False
    for src, stamp_file, stamp, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode != 0:
            failed = True
            sys.stderr.write("nvcc failed for %s:\n%s\n" % (src, out))
        else:
            if verbose or out.strip():
                sys.stderr.write(out)
            with open(stamp_file, "w") as f:
                f.write(stamp)
    if failed:
        raise RuntimeError("nvcc compilation failed")
    if procs or not os.path.exists(LIB):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart", "-ldl"]
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    variant = None
    extra = [a for a in sys.argv[1:] if a.startswith("-D")]
    for a in sys.argv[1:]:
        if a.startswith("--variant="):
            variant = a.split("=", 1)[1]
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=variant, extra_flags=extra))
