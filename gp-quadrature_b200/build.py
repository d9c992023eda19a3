"""Build libefgp_b200.so in-tree with nvcc for sm_100a (no torch extension machinery needed:
the boundary is a plain C ABI loaded with ctypes)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "efgp_b200")
LIB = os.path.join(OUT_DIR, "libefgp_b200.so")
OBJ_DIR = os.path.join(HERE, "build")

SOURCES = ["es_kernel.cpp", "pool.cpp", "timing.cpp", "points.cu", "spread.cu", "interp.cu", "nufft.cu", "toeplitz.cu", "cg.cu", "cg_small.cu", "lanczos.cu", "variance.cu", "fit.cu", "comm.cpp", "noalloc.cu", "capi.cu"]
HEADERS = ["common.cuh", "nufft.cuh", "toeplitz.cuh", "stream_ring.cuh", "fit.cuh", os.path.join("..", "..", "include", "efgp_b200.h")]

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-O3", "-x", "cu"]


def _newer(src_paths, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(p) > t for p in src_paths)


def build(force=False, verbose=True):
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    jobs = []
    objs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ_DIR, os.path.splitext(s)[0] + ".o")
        objs.append(obj)
        if force or _newer([src] + hdrs, obj):
            jobs.append([NVCC] + FLAGS + ["-c", src, "-o", obj])

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
        return r

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(run, jobs))
    if jobs or force or _newer(objs, LIB):
        run([NVCC, "-shared", "-o", LIB] + objs +
            ["-L/usr/local/cuda/lib64", "-lcufft", "-lcudart", "-ldl", "-Xlinker", "-rpath=/usr/local/cuda/lib64"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
