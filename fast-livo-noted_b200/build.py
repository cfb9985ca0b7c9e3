"""Builds libflivo_b200.so (sm_100a only) in-tree with nvcc.  No JIT cache, no torch extension machinery:
the product is a plain C-ABI shared library (include/flivo.h)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libflivo_b200.so")
SOURCES = ["map.cu", "fused.cu", "solve.cu", "stage.cu", "frontend.cu", "api.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: every float that feeds a comparison must round as separate mul/add (see csrc/flivo_dev.cuh)
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
         "-Xcompiler", "-fPIC,-fopenmp", "-I" + CSRC, "-I" + os.path.join(HERE, "..", "include")]


def _newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "flivo.h"),
                                                                os.path.abspath(__file__)]


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link libflivo_b200.so.  Returns the library path."""
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= _newest(_deps()):
        return LIB
    if not os.path.exists(NVCC):
        if os.path.exists(LIB):   # GPU box without a toolkit: use the prebuilt file that travelled with the repo
            return LIB
        raise RuntimeError("nvcc not found and no prebuilt libflivo_b200.so")
    os.makedirs(OBJ, exist_ok=True)

    def cc(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.check_call(cmd)
        return obj

    with ThreadPoolExecutor(max_workers=4) as ex:
        objs = list(ex.map(cc, SOURCES))
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-ldl", "-lgomp"]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return LIB


def build_host_test(force=False):
    """Compile the C++ caller of the drop-in classes (tests/cpp/test_dropin.cpp) against the library."""
    src = os.path.join(HERE, "..", "tests", "cpp", "test_dropin.cpp")
    out = os.path.join(OBJ, "test_dropin")
    deps = [src, LIB] + [os.path.join(HERE, "host", f) for f in os.listdir(os.path.join(HERE, "host"))]
    if not force and os.path.exists(out) and os.path.getmtime(out) >= _newest(deps):
        return out
    os.makedirs(OBJ, exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-I" + os.path.join(HERE, "host"), "-I" + os.path.join(HERE, "..", "include"),
                           src, "-o", out, "-L" + HERE, "-lflivo_b200", "-Wl,-rpath," + HERE, "-pthread"])
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
