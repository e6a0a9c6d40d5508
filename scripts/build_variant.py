#!/usr/bin/env python
"""A/B builds: compile the library with extra nvcc flags into gymothelloenv_b200/_lib_<name>/ (git-ignored).
usage: python scripts/build_variant.py <name> [-DFLAG ...] [--sizes 8,6]
On the GPU box an A/B run swaps the variant in:  cp gymothelloenv_b200/_lib_<name>/libothello_b200.so gymothelloenv_b200/_lib/
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import build as b  # noqa: E402

name = sys.argv[1]
flags = [a for a in sys.argv[2:] if a.startswith("-")]
b.LIB_DIR = os.path.join(b.PKG, "_lib_" + name)
b.OBJ_DIR = os.path.join(b.LIB_DIR, "obj")
b.LIB_PATH = os.path.join(b.LIB_DIR, "libothello_b200.so")
b.NVCC_FLAGS = b.NVCC_FLAGS + flags
print(b.build(force=True, verbose=True))
