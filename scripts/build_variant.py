#!/usr/bin/env python
"""A/B builds: compile the library with extra nvcc flags into gymothelloenv_b200/_lib_<name>/ (git-ignored).
usage: python scripts/build_variant.py <name> [-DFLAG ...] [--sizes 8,6]
--sizes: only these board sizes are recompiled with the flags; the objects of the other sizes are taken from the
main build (gymothelloenv_b200/_lib/obj, built first if stale).
On the GPU box an A/B run selects the variant with:  OTH_LIB_DIR=_lib_<name> python bench.py ...
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import build as b  # noqa: E402

name = sys.argv[1]
args = sys.argv[2:]
sizes = None
if "--sizes" in args:
    k = args.index("--sizes")
    sizes = [int(x) for x in args[k + 1].split(",")]
    del args[k:k + 2]
flags = [a for a in args if a.startswith("-")]
main_obj = b.OBJ_DIR
b.build()
b.LIB_DIR = os.path.join(b.PKG, "_lib_" + name)
b.OBJ_DIR = os.path.join(b.LIB_DIR, "obj")
b.LIB_PATH = os.path.join(b.LIB_DIR, "libothello_b200.so")
b.NVCC_FLAGS = b.NVCC_FLAGS + flags
print(b.build(force=True, verbose=True, only_sizes=sizes, reuse_obj_dir=main_obj))
import shutil  # noqa: E402
shutil.rmtree(b.OBJ_DIR)   # the snapshot sent to the GPU box is size-limited; the variant only needs its .so
