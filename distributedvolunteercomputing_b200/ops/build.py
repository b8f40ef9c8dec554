"""Build the native library in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m distributedvolunteercomputing_b200.ops.build [--force]

Output: distributedvolunteercomputing_b200/ops/libdvc_b200.so (git-ignored; it travels to
the GPU box with the gpurun snapshot).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
LIB_PATH = os.path.join(HERE, "libdvc_b200.so")
STAMP = os.path.join(HERE, ".libdvc_b200.stamp")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
              "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest() -> str:
    h = hashlib.sha256()
    files = sources() + [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cuh")]
    files += [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE))]
    for f in files:
        h.update(f.encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def nvcc_path() -> str:
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build(force: bool = False, verbose: bool = False) -> str:
    dig = _digest()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(STAMP) and open(STAMP).read().strip() == dig:
        return LIB_PATH
    objdir = os.path.join(ROOT, "build", "obj")
    os.makedirs(objdir, exist_ok=True)
    nvcc = nvcc_path()
    procs = []
    objs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, "-I", INCLUDE, "-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {os.path.basename(src)}\n{out}")
        if p.returncode != 0:
            failed = True
    with open(os.path.join(ROOT, "build", "nvcc.log"), "w") as f:
        f.write("\n".join(log))
    if failed:
        sys.stderr.write("\n".join(log))
        raise RuntimeError("nvcc failed; see build/nvcc.log")
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH, *objs]
    subprocess.run(cmd, check=True)
    with open(STAMP, "w") as f:
        f.write(dig)
    if verbose:
        print("\n".join(log))
    return LIB_PATH


def build_tsan(force: bool = False) -> str:
    """ThreadSanitizer build of the library's HOST code plus the harness tests/native/tsan_host.cpp (no GPU needed: the
    harness drives every entry point up to the first driver call).  Everything lands under build/tsan/ - the product
    library is untouched.  Returns the path of the harness executable."""
    out = os.path.join(ROOT, "build", "tsan")
    os.makedirs(out, exist_ok=True)
    lib = os.path.join(out, "libdvc_b200_tsan.so")
    exe = os.path.join(out, "tsan_host")
    stamp = os.path.join(out, "stamp")
    harness = os.path.join(ROOT, "tests", "native", "tsan_host.cpp")
    h = hashlib.sha256((_digest() + open(harness).read()).encode()).hexdigest()
    if not force and os.path.exists(exe) and os.path.exists(stamp) and open(stamp).read().strip() == h:
        return exe
    nvcc = nvcc_path()
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-g", "-std=c++17", "--expt-relaxed-constexpr",
             "-Xcompiler", "-fPIC,-fsanitize=thread,-fno-omit-frame-pointer"]
    procs, objs = [], []
    for src in sources():
        obj = os.path.join(out, os.path.basename(src) + ".o")
        objs.append(obj)
        procs.append((src, subprocess.Popen([nvcc, *flags, "-I", INCLUDE, "-c", src, "-o", obj], stdout=subprocess.PIPE,
                                            stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        text, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"tsan build of {os.path.basename(src)} failed:\n{text}")
    subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fsanitize=thread", "-o", lib, *objs],
                   check=True)
    subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fno-omit-frame-pointer", "-I", INCLUDE, harness,
                    "-o", exe, "-L", out, "-ldvc_b200_tsan", f"-Wl,-rpath,{out}", "-lpthread"], check=True)
    with open(stamp, "w") as f:
        f.write(h)
    return exe


if __name__ == "__main__":
    if "--tsan" in sys.argv:
        exe = build_tsan(force="--force" in sys.argv)
        sys.exit(subprocess.run([exe], env={**os.environ, "TSAN_OPTIONS": "halt_on_error=0 exitcode=66"}).returncode)
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
