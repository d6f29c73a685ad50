"""Build the in-tree native library ``fed_tgan_b200/_native/libfedtgan_b200.so``.

Plain ``nvcc`` (sm_100a only, ``-lineinfo`` so ncu source pages map to ``csrc/``), one
object per translation unit compiled in parallel, linked into a single C-ABI shared
object that ``ops/native.py`` loads with ``ctypes``.  The ``.so`` is git-ignored but
travels to the GPU box with the repo snapshot.
"""

from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "fed_tgan_b200", "_native")
LIB = os.path.join(OUT_DIR, "libfedtgan_b200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "--use_fast_math=false"]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def sources():
    cu = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    cpp = sorted(f for f in os.listdir(CSRC) if f.endswith(".cpp"))
    return cu, cpp


HASH_FILE = LIB + ".srchash"


def _source_hash():
    """SHA-256 over the contents of every source that goes into the library (names included)."""
    import hashlib

    h = hashlib.sha256()
    deps = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [os.path.join(ROOT, "include", "fedtgan_b200.h")]
    for d in deps:
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(ARCH + NVCC_FLAGS).encode())
    return h.hexdigest()


def _stale():
    """The shipped library is reused iff it was built from exactly these sources.  (Contents, not mtimes: copying
    the repo to another machine must not trigger a 1-2 minute rebuild in every test run.)"""
    if not os.path.exists(LIB) or not os.path.exists(HASH_FILE):
        return True
    try:
        return open(HASH_FILE).read().strip() != _source_hash()
    except OSError:
        return True


def build(force=False, verbose=False):
    """Compile (if stale) and return the library path."""
    if not force and not _stale():
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    # One builder at a time: under torchrun every rank imports ops.native at once, and a snapshot copy of the
    # repo can make the shipped .so look older than the sources.  The first rank builds, the others wait for the
    # lock and then find a fresh library.
    import fcntl

    lock = open(os.path.join(OUT_DIR, ".build.lock"), "w")
    fcntl.flock(lock, fcntl.LOCK_EX)
    try:
        if not force and not _stale():
            return LIB
        return _build_locked(verbose)
    finally:
        fcntl.flock(lock, fcntl.LOCK_UN)
        lock.close()


def _build_locked(verbose=False):
    obj_dir = os.path.join(OUT_DIR, "obj")
    os.makedirs(obj_dir, exist_ok=True)
    cu, cpp = sources()
    flags = [f for f in NVCC_FLAGS if not f.startswith("--use_fast_math")]

    def compile_one(name):
        src = os.path.join(CSRC, name)
        obj = os.path.join(obj_dir, name + ".o")
        if name.endswith(".cu"):
            cmd = [_nvcc(), *ARCH, *flags, "-Xptxas", "-v", "-c", src, "-o", obj]
        else:
            cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-c", src, "-o", obj]
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"compile failed: {' '.join(cmd)}\n{res.stdout}")
        with open(obj + ".log", "w") as fh:
            fh.write(res.stdout)
        if verbose:
            print(res.stdout)
        return obj

    with ThreadPoolExecutor(max_workers=8) as pool:
        objs = list(pool.map(compile_one, cu + cpp))
    cmd = [_nvcc(), *ARCH, "-shared", "-o", LIB + ".tmp", *objs, "-lpthread"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed: {res.stdout}")
    os.replace(LIB + ".tmp", LIB)
    with open(HASH_FILE + ".tmp", "w") as fh:
        fh.write(_source_hash())
    os.replace(HASH_FILE + ".tmp", HASH_FILE)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
