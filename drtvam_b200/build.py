"""In-tree build of the CUDA C-ABI library for sm_100a (B200).

``python -m drtvam_b200.build`` compiles ``csrc/*.cu`` with nvcc into
``drtvam_b200/lib/libdrtvam_b200.so``.  The ``.so`` is git-ignored but travels
to the GPU box with the repo snapshot.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libdrtvam_b200.so")
MICRO = os.path.join(LIBDIR, "libdrtvam_microbench.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-shared",
    "-Xptxas", "-v",
]

LIB_SOURCES = ["abi.cu", "general_kernels.cu", "planar_kernels.cu", "planarz_kernels.cu", "planarj_kernels.cu",
               "loss_kernels.cu", "lbfgs_kernels.cu",
               "mesh_kernels.cu", "inside_kernels.cu", "filter_kernels.cu"]
MICRO_SOURCES = ["microbench.cu"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA toolkit is required to build drtvam_b200")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _build_one(out: str, sources, force: bool, verbose: bool) -> bool:
    srcs = [os.path.join(CSRC, s) for s in sources if os.path.exists(os.path.join(CSRC, s))]
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(HERE, "..", "include", "drtvam_b200.h")]
    stamp = out + ".sha256"
    digest = _digest([d for d in deps if os.path.isfile(d)])
    if not force and os.path.exists(out) and os.path.exists(stamp):
        with open(stamp) as f:
            if f.read().strip() == digest:
                return False
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    # the image exports CC=/opt/gcc/bin/gcc; nvcc needs the system host compiler
    ccbin = ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    compile_flags = [f for f in NVCC_FLAGS if f != "-shared"]

    def compile_one(src):
        obj = os.path.join(objdir, os.path.basename(out) + "." + os.path.basename(src) + ".o")
        cmd = [_nvcc(), *ccbin, *compile_flags, "-c", "-o", obj, src]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        return obj, " ".join(cmd) + "\n" + proc.stdout + proc.stderr, proc.returncode

    # one translation unit per nvcc process, all of them at once (nvcc cross-compiles without a GPU)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 4)) as pool:
        results = list(pool.map(compile_one, srcs))
    log = "".join(r[1] for r in results)
    rc = max(r[2] for r in results)
    if rc == 0:
        cmd = [_nvcc(), *ccbin, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
               "-o", out, *[r[0] for r in results], "-ldl"]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        log += " ".join(cmd) + "\n" + proc.stdout + proc.stderr
        rc = proc.returncode
    with open(out + ".buildlog", "w") as f:
        f.write(log)
    if rc != 0:
        sys.stderr.write(log)
        raise RuntimeError(f"nvcc failed building {os.path.basename(out)}")
    if verbose:
        print(log)
    with open(stamp, "w") as f:
        f.write(digest)
    return True


def build(force: bool = False, verbose: bool = False) -> str:
    _build_one(LIB, LIB_SOURCES, force, verbose)
    if all(os.path.exists(os.path.join(CSRC, s)) for s in MICRO_SOURCES):
        _build_one(MICRO, MICRO_SOURCES, force, verbose)
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
