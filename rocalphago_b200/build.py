"""Builds librocalphago_b200.so (hand-written sm_100a CUDA + the C ABI) with nvcc, in-tree."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "librocalphago_b200.so")
SOURCES = ["features.cu", "conv.cu"]
HEADERS = ["go_bb.cuh", "rag_common.h", "../../include/rocalphago_b200.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    files = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.exists(f) and os.path.getmtime(f) > t for f in files)


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
        obj = os.path.join(LIB_DIR, src.replace(".cu", ".o"))
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", path, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.check_call(cmd)
        objs.append(obj)
    subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH]
                          + objs + ["-lcuda", "-lpthread"])
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
