"""Compile the CUDA library in-tree for sm_100a (`python -m sparc_b200.build`)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.environ.get("SPARC_LIB_PATH") or os.path.join(LIB_DIR, "libsparc_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # fp64 scalar path must round like Python floats; the stencil uses explicit FMAs
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """One `nvcc -c` per translation unit, in parallel, then one link into libsparc_b200.so."""
    if not force and not _stale():
        return LIB_PATH
    from concurrent.futures import ThreadPoolExecutor

    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    obj_dir = LIB_PATH + ".obj"
    os.makedirs(obj_dir, exist_ok=True)
    extra = os.environ.get("SPARC_NVCC_EXTRA", "").split()     # experiments only (e.g. -DSPARC_MINB_SMALL=5)

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc] + NVCC_FLAGS + extra + ["-I", INCLUDE, "-c", "-o", obj, src]
        res = subprocess.run(cmd, capture_output=True, text=True)
        return obj, " ".join(cmd) + "\n" + res.stdout + res.stderr, res.returncode

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        results = list(pool.map(compile_one, sources()))
    log = "".join(r[1] for r in results)
    ok = all(r[2] == 0 for r in results)
    if ok:
        cmd = [nvcc, "-shared", "-o", LIB_PATH] + [r[0] for r in results]
        res = subprocess.run(cmd, capture_output=True, text=True)
        log += " ".join(cmd) + "\n" + res.stdout + res.stderr
        ok = res.returncode == 0
    with open(LIB_PATH + ".build.log", "w") as f:
        f.write(log)
    if not ok:
        raise RuntimeError("nvcc failed:\n" + log[-4000:])
    if verbose:
        print(log)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
