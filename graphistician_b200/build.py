"""Build libgk.so (hand-written sm_100a CUDA behind the C ABI of include/gk.h) in-tree with nvcc.

Every csrc/*.cu is compiled to an object file (in parallel, objects under csrc/_obj/) and linked into
graphistician_b200/libgk.so; only the translation units whose sources changed are recompiled."""
import glob
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libgk.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "gk.h")]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in sources() + _headers())


def _obj_of(src):
    return os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")


def _stale(src):
    o = _obj_of(src)
    if not os.path.exists(o):
        return True
    t = os.path.getmtime(o)
    return any(os.path.getmtime(d) > t for d in [src] + _headers())


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo ... csrc/*.cu -> graphistician_b200/libgk.so"""
    if not force and not needs_build():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(OBJ, exist_ok=True)
    todo = [s for s in sources() if force or _stale(s)]

    def compile_one(src):
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", _obj_of(src), src]
        subprocess.check_call(cmd, cwd=CSRC)

    with ThreadPoolExecutor(max_workers=max(1, min(len(todo), os.cpu_count() or 1))) as ex:
        list(ex.map(compile_one, todo))
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] +
                          [_obj_of(s) for s in sources()], cwd=CSRC)
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="-f" in sys.argv or "-v" in sys.argv, verbose="-v" in sys.argv))
