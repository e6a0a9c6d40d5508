"""Builds libothello_b200.so (sm_100a) in-tree with nvcc -- no torch involved.

    python -m gymothelloenv_b200.build [--force]

One translation unit per board size (csrc/oth_size.cu with -DOTH_S=n) compiled in
parallel, plus the C ABI layer (csrc/oth_cabi.cu); linked into
gymothelloenv_b200/_lib/libothello_b200.so.  The .so is git-ignored but travels to the
GPU box with the repo snapshot.
"""
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
INCLUDE = os.path.join(os.path.dirname(PKG), "include")
LIB_DIR = os.path.join(PKG, "_lib")
OBJ_DIR = os.path.join(PKG, "_lib", "obj")
LIB_PATH = os.path.join(LIB_DIR, "libothello_b200.so")
BOARD_SIZES = list(range(4, 17))   # must match OTH_DECL list in oth_cabi.cu / OTH_MAX_BOARD_SIZE

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-diag-suppress", "177"]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libothello_b200.so")
    return exe


def _sources():
    """Every file the library is compiled from: all of csrc/ and include/ (a fixed list once missed a new
    header, so edits to it did not trigger a rebuild)."""
    files = [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if n.endswith((".cu", ".cuh", ".h"))]
    files += [os.path.join(INCLUDE, n) for n in sorted(os.listdir(INCLUDE)) if n.endswith(".h")]
    return files


def source_digest():
    h = hashlib.sha256()
    for p in _sources():
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(str(BOARD_SIZES).encode())
    return h.hexdigest()


def is_fresh():
    stamp = LIB_PATH + ".digest"
    return os.path.isfile(LIB_PATH) and os.path.isfile(stamp) and open(stamp).read().strip() == source_digest()


def _compile(args):
    src, obj, defs = args
    cmd = [_nvcc()] + NVCC_FLAGS + defs + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return obj


def build(force=False, verbose=False, only_sizes=None, reuse_obj_dir=None):
    """only_sizes / reuse_obj_dir: A/B variant builds (scripts/build_variant.py) recompile a few sizes and link the
    other sizes' objects from an existing build."""
    if not force and is_fresh():
        return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    jobs = [(os.path.join(CSRC, "oth_size.cu"), os.path.join(OBJ_DIR, "oth_size_%d.o" % s), ["-DOTH_S=%d" % s])
            for s in BOARD_SIZES]
    jobs.append((os.path.join(CSRC, "oth_cabi.cu"), os.path.join(OBJ_DIR, "oth_cabi.o"), []))
    reused = []
    if only_sizes is not None:
        keep = []
        for j in jobs:
            size = int(j[2][0].split("=")[1]) if j[2] else None
            if size is None or size in only_sizes:
                keep.append(j)
            else:
                shutil.copyfile(os.path.join(reuse_obj_dir, os.path.basename(j[1])), j[1])
                reused.append(j[1])
        jobs = keep
    # big boards first: they take the longest
    jobs.sort(key=lambda j: -int(j[2][0].split("=")[1]) if j[2] else 0)
    with cf.ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 4))) as ex:
        objs = list(ex.map(_compile, jobs)) + reused
    tmp = LIB_PATH + ".tmp%d" % os.getpid()
    cmd = [_nvcc(), "-shared", "-o", tmp] + sorted(objs) + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stdout))
    os.replace(tmp, LIB_PATH)
    with open(LIB_PATH + ".digest", "w") as f:
        f.write(source_digest())
    if verbose:
        print("built", LIB_PATH)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
