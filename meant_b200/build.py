"""Build libmeant_b200.so (sm_100a only) in-tree with nvcc.

    python -m meant_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the repo snapshot.  There is no JIT
and no fallback: `meant_b200._native` raises if the library is missing.
"""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "_lib")
LIB = os.path.join(LIBDIR, "libmeant_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--use_fast_math=false", "-Xptxas", "-v"]
FLAGS.remove("--use_fast_math=false")  # exact erf/exp/div: parity first
if os.environ.get("MEANT_PROF"):         # in-kernel phase timers (tools/kbench.py --prof); never set for product builds
    FLAGS.append("-DMEANT_PROF")


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest():
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)) + ["../../include/meant_b200.h"]:
        p = os.path.join(CSRC, f)
        if os.path.isfile(p):
            h.update(f.encode())
            h.update(open(p, "rb").read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _compile(src):
    obj = os.path.join(LIBDIR, src[:-3] + ".o")
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(os.path.join(LIBDIR, src[:-3] + ".ptxas.log"), "w") as f:
        f.write(r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stderr[-4000:]}")
    return obj


def build(force=False, verbose=True):
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, "build.sha256")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        if verbose:
            print("meant_b200: library up to date", LIB)
        return LIB
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(_compile, sources()))
    r = subprocess.run([NVCC, "-shared", "-o", LIB, *objs], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stderr)
    open(stamp, "w").write(dig)
    if verbose:
        print("meant_b200: built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
