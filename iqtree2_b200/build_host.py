"""In-tree build of the C++ host layer (iqtree2_b200/host/ -> libb200host.so).  It resolves the kernel C ABI from
libb200lk.so at run time (dlopen), so it compiles and loads without CUDA."""
import glob
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libb200host.so")
# the same sources with -DB200_TESTING: adds b200host_set_kernel_library, the hook tests/ use to run the host logic over
# the CPU oracle where no GPU exists.  Loaded by tests only (iqtree2_b200.host.set_kernel_library).
TEST_LIB = os.path.join(HERE, "libb200host_testing.so")
CXX = os.environ.get("B200_CXX", "/usr/bin/g++")  # the system compiler: its libstdc++.so is the one python extensions share
FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-Wall", "-Wno-unused-function"]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "host", "*.cpp")))


def needs_build(lib=LIB):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = sources() + glob.glob(os.path.join(HERE, "host", "*.h")) + [os.path.join(HERE, "..", "include", "b200lk.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, testing=False):
    lib = TEST_LIB if testing else LIB
    if not force and not needs_build(lib):
        return lib
    cmd = [CXX] + FLAGS + (["-DB200_TESTING"] if testing else []) + ["-o", lib] + sources() + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return lib


if __name__ == "__main__":
    build(force=True, verbose=True)
    build(force=True, verbose=True, testing=True)
