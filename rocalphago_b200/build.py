"""Builds librocalphago_b200.so (hand-written sm_100a CUDA + the C ABI) with nvcc, in-tree."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "librocalphago_b200.so")
SOURCES = ["features.cu", "conv.cu", "tree.cpp"]
HEADERS = ["go_bb.cuh", "rag_common.h", "../../include/rocalphago_b200.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


STAMP_PATH = os.path.join(LIB_DIR, "sources.sha1")


def _source_digest():
    """Content hash of everything the library is built from (modification times do not survive a copy
    of the tree to another machine; contents do)."""
    import hashlib
    h = hashlib.sha1(" ".join(NVCC_FLAGS).encode())
    for rel in SOURCES + HEADERS:
        path = os.path.join(CSRC, rel)
        if os.path.exists(path):
            with open(path, "rb") as f:
                h.update(rel.encode() + b"\0" + f.read())
    return h.hexdigest()


def _stale():
    """True if the shared library is missing or was built from other sources than the ones in the tree."""
    if not os.path.exists(LIB_PATH) or not os.path.exists(STAMP_PATH):
        return True
    with open(STAMP_PATH) as f:
        return f.read().strip() != _source_digest()


def build_library(force=False, verbose=False):
    """Compiles every .cu under csrc/ for sm_100a and links them into one shared library."""
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    objs = []
    for src in SOURCES:
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(LIB_DIR, os.path.splitext(src)[0] + ".o")
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", path, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.check_call(cmd)
        objs.append(obj)
    subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH]
                          + objs + ["-lcuda", "-lpthread"])
    with open(STAMP_PATH, "w") as f:
        f.write(_source_digest() + "\n")
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
