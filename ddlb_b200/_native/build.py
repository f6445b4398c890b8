"""In-tree build of ``libddlb_b200.so`` for sm_100a (cross-compiles without a GPU).

    python -m ddlb_b200._native.build [--force] [--verbose]
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
LIB_PATH = os.path.join(HERE, "libddlb_b200.so")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC or add /usr/local/cuda/bin to PATH)")


def sources() -> List[str]:
    out = []
    for base in (CSRC, INCLUDE):
        for name in sorted(os.listdir(base)):
            if name.endswith((".cu", ".cuh", ".inl", ".h")):
                out.append(os.path.join(base, name))
    return out


HASH_PATH = LIB_PATH + ".srchash"


def source_hash() -> str:
    import hashlib
    h = hashlib.sha256()
    for path in sources():
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def is_stale() -> bool:
    """True when the library is missing or was built from different sources. Compares a
    content hash stored beside the library, not mtimes: a snapshot copy of the tree may
    change mtimes, and a spurious rebuild in every rank of a multi-GPU run is not wanted."""
    if not os.path.exists(LIB_PATH):
        return True
    try:
        with open(HASH_PATH) as fh:
            return fh.read().strip() != source_hash()
    except OSError:
        return True


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    digest = source_hash()
    tmp = f"{LIB_PATH}.{os.getpid()}.tmp"          # concurrent builders never share a file
    cmd = [nvcc_path(), *ARCH_FLAGS, "-O3", "-lineinfo", "-std=c++17", "-shared",
           "-Xcompiler", "-fPIC", "-I", INCLUDE, os.path.join(CSRC, "ddlb_b200.cu"),
           "-o", tmp]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd))
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        try:
            os.remove(tmp)
        except OSError:
            pass
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    os.replace(tmp, LIB_PATH)                      # atomic: loaders see the old or the new file
    with open(f"{HASH_PATH}.{os.getpid()}.tmp", "w") as fh:
        fh.write(digest + "\n")
    os.replace(f"{HASH_PATH}.{os.getpid()}.tmp", HASH_PATH)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
