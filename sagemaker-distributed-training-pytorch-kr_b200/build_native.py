"""In-tree build of the native library (sm_100a only).

``python -m b200dist.build_native`` (or ``__graft_entry__.build()``) compiles every ``csrc/*.cu``
with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` and every ``csrc/host/*.cpp`` with
g++, then links ``lib/libb200dist.so`` (CUDA kernels + C ABI) and ``lib/libb200dist_host.so`` (CPU-side
runtime: DDP bucket planner, dataset index builders, pipeline schedule generator).  nvcc cross-compiles
without a GPU.  Objects are cached by source hash under ``build/`` so rebuilds are incremental.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(PKG_DIR)
CSRC = os.path.join(REPO, "csrc")
INCLUDE = os.path.join(REPO, "include")
BUILD = os.path.join(REPO, "build")
LIB_DIR = os.path.join(PKG_DIR, "lib")
CUDA_LIB = os.path.join(LIB_DIR, "libb200dist.so")
HOST_LIB = os.path.join(LIB_DIR, "libb200dist_host.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-I" + INCLUDE, "-I" + CSRC,
]
CXX_FLAGS = ["-std=c++17", "-O3", "-fPIC", "-march=x86-64-v3", "-I" + INCLUDE, "-I" + CSRC]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths, extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def _headers():
    hs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cuh", ".h"))]
    hs += [os.path.join(INCLUDE, f) for f in sorted(os.listdir(INCLUDE)) if f.endswith(".h")]
    return hs


def _compile(cmd, obj, stamp, tag, verbose):
    stamp_file = obj + ".stamp"
    if os.path.exists(obj) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return obj
    if verbose:
        print("[build]", tag, flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    with open(stamp_file, "w") as f:
        f.write(stamp)
    return obj


def build(verbose: bool = True, force: bool = False) -> None:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    nvcc = _nvcc()
    hdrs = _headers()
    cu = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
    host_dir = os.path.join(CSRC, "host")
    cpp = [os.path.join(host_dir, f) for f in sorted(os.listdir(host_dir)) if f.endswith(".cpp")] if os.path.isdir(host_dir) else []
    if force:
        for f in os.listdir(BUILD):
            if f.endswith(".stamp"):
                os.remove(os.path.join(BUILD, f))

    jobs = []
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        for src in cu:
            obj = os.path.join(BUILD, os.path.basename(src) + ".o")
            stamp = _digest([src] + hdrs, " ".join(NVCC_FLAGS))
            jobs.append(("cu", ex.submit(_compile, [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj], obj, stamp, os.path.basename(src), verbose)))
        for src in cpp:
            obj = os.path.join(BUILD, "host_" + os.path.basename(src) + ".o")
            stamp = _digest([src] + hdrs, " ".join(CXX_FLAGS))
            jobs.append(("cpp", ex.submit(_compile, ["g++"] + CXX_FLAGS + ["-c", src, "-o", obj], obj, stamp, os.path.basename(src), verbose)))
        cu_objs = [j.result() for k, j in jobs if k == "cu"]
        cpp_objs = [j.result() for k, j in jobs if k == "cpp"]

    def _link(cmd, out, objs):
        newest = max(os.path.getmtime(o) for o in objs)
        if os.path.exists(out) and os.path.getmtime(out) >= newest and not force:
            return
        if verbose:
            print("[link]", os.path.basename(out), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")

    if cu_objs:
        _link([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", CUDA_LIB] + cu_objs, CUDA_LIB, cu_objs)
    if cpp_objs:
        _link(["g++", "-shared", "-o", HOST_LIB] + cpp_objs + ["-lpthread"], HOST_LIB, cpp_objs)


if __name__ == "__main__":
    build(verbose=True, force="--force" in sys.argv)
    print("ok:", CUDA_LIB, HOST_LIB)
