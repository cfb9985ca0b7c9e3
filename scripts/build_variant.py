#!/usr/bin/env python
"""A/B builds of libflivo_b200.so: python scripts/build_variant.py <name> [-DFLAG=..]...  -> fast-livo-noted_b200/variants/<name>.so
(loaded through FLIVO_LIB=variants/<name>.so, see fast-livo-noted_b200/__init__.py; variants/ is not tracked)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.load_package()
from importlib import import_module  # noqa: E402
B = import_module(ge.PKG_NAME + ".build")
name, extra = sys.argv[1], sys.argv[2:]
# --src DIR: compile a (patched) copy of csrc/ instead of the tree's, so that an experiment never leaves the tree half-edited
SRC = B.CSRC
if "--src" in extra:
    i = extra.index("--src"); SRC = os.path.abspath(extra[i + 1]); extra = extra[:i] + extra[i + 2:]
FLAGS = [("-I" + SRC) if fl == "-I" + B.CSRC else fl for fl in B.FLAGS]
out_dir = os.path.join(B.HERE, "variants")
obj_dir = os.path.join(B.OBJ, "variant_" + name)
os.makedirs(out_dir, exist_ok=True)
os.makedirs(obj_dir, exist_ok=True)


def cc(src):
    obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
    subprocess.check_call([B.NVCC] + FLAGS + extra + ["-c", os.path.join(SRC, src), "-o", obj])
    return obj


with ThreadPoolExecutor(max_workers=5) as ex:
    objs = list(ex.map(cc, B.SOURCES))
lib = os.path.join(out_dir, name + ".so")
subprocess.check_call([B.NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs + ["-ldl", "-lgomp"])
print(lib)
