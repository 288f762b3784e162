"""In-tree build of the C++ host layer (iqtree2_b200/host/ -> libb200host.so).  It resolves the kernel C ABI from
libb200lk.so at run time (dlopen), so it compiles and loads without CUDA."""
import glob
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libb200host.so")
CXX = os.environ.get("B200_CXX", "/usr/bin/g++")  # the system compiler: its libstdc++.so is the one python extensions share
FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-Wall", "-Wno-unused-function"]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "host", "*.cpp")))


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(HERE, "host", "*.h")) + [os.path.join(HERE, "..", "include", "b200lk.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    cmd = [CXX] + FLAGS + ["-o", LIB] + sources() + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    build(force=True, verbose=True)
