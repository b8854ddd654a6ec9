"""Build libtracer_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

Every .cu is compiled to its own object (in parallel, re-used while it and the headers are unchanged) and the
objects are linked into the shared library."""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, "libtracer_b200.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = ["tracer_b200.cu", "nearest.cu", "bvh_build.cu", "sort.cu", "bounce.cu", "sources.cu", "trace.cu", "shadow_grid.cu", "probe.cu"]
HEADERS = ["common.cuh", "geom.cuh", "optics.cuh", "philox.cuh",
           os.path.join("..", "..", "include", "tracer_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _digest(paths, extra):
    h = hashlib.sha1(" ".join(extra).encode())
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def build(force=False, verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("TRC_NVCC_EXTRA", "").split()      # tuning builds: -DTRC_WALK_WARPS=6 ...
    flags = NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else [])
    headers = [os.path.join(HERE, h) for h in HEADERS]
    os.makedirs(OBJ_DIR, exist_ok=True)
    jobs, objs = [], []
    for src in SOURCES:
        path = os.path.join(HERE, src)
        tag = _digest([path] + headers + [os.path.abspath(__file__)], flags)
        obj = os.path.join(OBJ_DIR, "%s.%s.o" % (os.path.splitext(src)[0], tag))
        objs.append(obj)
        if force or verbose or not os.path.exists(obj):
            for old in os.listdir(OBJ_DIR):                     # drop stale objects of this source
                if old.startswith(os.path.splitext(src)[0] + ".") and os.path.join(OBJ_DIR, old) != obj:
                    os.remove(os.path.join(OBJ_DIR, old))
            jobs.append([nvcc] + flags + ["-c", path, "-o", obj])
    if not jobs and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(o) for o in objs):
        return LIB

    def run(cmd):
        return cmd, subprocess.run(cmd, capture_output=True, text=True)
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as pool:
        for cmd, res in pool.map(run, jobs):
            if res.returncode != 0:
                sys.stderr.write(res.stdout + res.stderr)
                raise RuntimeError("nvcc failed: " + " ".join(cmd))
            if verbose:
                sys.stderr.write(res.stdout + res.stderr)
    res = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", LIB],
                         capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libtracer_b200.so")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
