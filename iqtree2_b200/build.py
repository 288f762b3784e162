"""In-tree build of the CUDA library (sm_100a only; nvcc cross-compiles without a GPU).
One object per translation unit (compiled in parallel, rebuilt only when stale), then one link."""
import glob
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libb200lk.so")
OBJDIR = os.path.join(HERE, "csrc", "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CFLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
          "-Xcompiler", "-fPIC,-fopenmp"]  # OpenMP: the per-shard host fan-out of shard.cu
LFLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC,-fopenmp"]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))


def _headers():
    return glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + [os.path.join(HERE, "..", "include", "b200lk.h")]


def _obj(src):
    return os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")


def _stale(src, force):
    o = _obj(src)
    if force or not os.path.exists(o):
        return True
    t = os.path.getmtime(o)
    return any(os.path.getmtime(d) > t for d in [src] + _headers())


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in sources() + _headers())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJDIR, exist_ok=True)
    todo = [s for s in sources() if _stale(s, force)]

    def cc(src):
        cmd = [NVCC] + CFLAGS + ["-c", "-o", _obj(src), src]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as ex:
        list(ex.map(cc, todo))
    cmd = [NVCC] + LFLAGS + ["-o", LIB] + [_obj(s) for s in sources()]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    import sys
    build(force="--force" in sys.argv, verbose=True)
