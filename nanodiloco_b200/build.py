"""Builds libnanodiloco_b200.so in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU, so this runs in the CPU-only build container;
the built library travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libnanodiloco_b200.so")
SOURCES = ["nd_kernels.cu", "nd_sync.cu", "nd_peer.cu"]
HEADERS = ["nd_common.cuh", "nd_error.h", "nd_tma.cuh", os.path.join("..", "..", "include", "nanodiloco_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "shared",
    "-Xlinker", "-rpath,/usr/local/cuda/lib64",
]


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; libnanodiloco_b200.so cannot be built")
    return nvcc


def is_stale() -> bool:
    """True when a source file is newer than the built library (a developer convenience; file
    times do not survive every copy of the tree, so nothing rebuilds on this signal alone)."""
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def needs_build() -> bool:
    """The library is built when it is MISSING (or when ND_REBUILD=1 asks for it).  An existing
    library is trusted: with one process per GPU, eight ranks deciding from file times to rebuild
    the same file at once is how a rank ends up loading a half-written library."""
    return not os.path.exists(SO) or os.environ.get("ND_REBUILD") == "1"


def build(force: bool = False, verbose: bool = False) -> str:
    """Compiles into a temporary file and renames it into place under a file lock, so concurrent
    callers (ranks of one job) either build once or wait, and nobody ever sees a partial file."""
    import fcntl
    if not force and not needs_build():
        return SO
    lock_path = SO + ".lock"
    with open(lock_path, "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and os.path.exists(SO) and os.environ.get("ND_REBUILD") != "1":
                return SO            # another process built it while we waited
            tmp = f"{SO}.tmp.{os.getpid()}"
            cmd = [_nvcc()] + NVCC_FLAGS + [os.path.join(CSRC, f) for f in SOURCES] + ["-o", tmp, "-ldl"]
            if verbose:
                cmd.insert(1, "-Xptxas")
                cmd.insert(2, "-v")
                print(" ".join(cmd))
            try:
                subprocess.check_call(cmd)
                os.replace(tmp, SO)
            finally:
                if os.path.exists(tmp):
                    os.remove(tmp)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return SO


if __name__ == "__main__":
    import sys
    # developer entry point: rebuild when asked to or when a source file is newer than the library
    print(build(force="--force" in sys.argv or is_stale(), verbose="-v" in sys.argv))
