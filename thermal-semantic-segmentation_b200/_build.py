"""In-tree nvcc build of libtss_b200.so (sm_100a only, no JIT cache, no multi-arch).

``python -m b200seg._build`` or ``b200seg._build.build()``; ``__graft_entry__.build()`` calls it.
The shared object lands beside this file so that it travels with the repo snapshot to the GPU box.
"""
from __future__ import annotations

import fcntl
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ_DIR = os.path.join(HERE, "build") if not os.environ.get("TSS_LIB") else os.environ["TSS_LIB"] + ".build"
# TSS_LIB (tuning experiments only): build / load a variant library, compiled with the extra flags in TSS_NVCC_EXTRA
LIB_PATH = os.environ.get("TSS_LIB") or os.path.join(HERE, "libtss_b200.so")
STAMP = os.path.join(OBJ_DIR, "sources.sha256")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
] + os.environ.get("TSS_NVCC_EXTRA", "").split()


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libtss_b200.so cannot be built (there is no CPU fallback)")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest() -> str:
    h = hashlib.sha256()
    files = sources() + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    files.append(os.path.join(os.path.dirname(HERE), "include", "tss_b200.h"))
    for f in files:
        h.update(os.path.basename(f).encode())      # names, not absolute paths: the digest must survive a move of the tree
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def source_digest() -> str:
    """sha256 over csrc/, the header and the nvcc flags; compiled into the library (tss_source_digest)."""
    return _digest()


def is_current() -> bool:
    try:
        with open(STAMP) as fh:
            return os.path.exists(LIB_PATH) and fh.read().strip() == _digest()
    except OSError:
        return False


class _BuildLock:
    """Exclusive inter-process lock (one rank builds, the others wait and then find a current library)."""

    def __enter__(self):
        os.makedirs(OBJ_DIR, exist_ok=True)
        self.fh = open(os.path.join(OBJ_DIR, ".lock"), "w")
        fcntl.flock(self.fh, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        fcntl.flock(self.fh, fcntl.LOCK_UN)
        self.fh.close()


def embedded_digest(path: str = None) -> str:
    """The source digest compiled into an existing library ('' when it has none).  Read from the file, not through
    dlopen: loading a stale library would keep it mapped under this path for the rest of the process."""
    marker = b"TSS_SOURCE_DIGEST="
    try:
        with open(path or LIB_PATH, "rb") as fh:
            blob = fh.read()
    except OSError:
        return ""
    i = blob.find(marker)
    if i < 0:
        return ""
    j = blob.find(b"\0", i)
    return blob[i + len(marker):j].decode(errors="replace")


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a and link libtss_b200.so.  Returns the library path.
    Safe under torchrun: the build runs under a file lock with per-process temporary names."""
    if not force and is_current():
        return LIB_PATH
    with _BuildLock():
        if not force and is_current():      # another rank built it while this one waited
            return LIB_PATH
        return _build_locked(verbose)


def ensure_current() -> str:
    """The library path, after making sure the library was compiled from the sources in csrc/ (rebuilding
    under the lock when it was not).  Raises RuntimeError when that needs nvcc and there is none."""
    want = _digest()
    if os.path.exists(LIB_PATH) and embedded_digest() == want:
        return LIB_PATH
    with _BuildLock():
        if os.path.exists(LIB_PATH) and embedded_digest() == want:   # another rank rebuilt it meanwhile
            return LIB_PATH
        return _build_locked(False)


def _build_locked(verbose: bool) -> str:
    nvcc = _nvcc()
    digest = _digest()
    pid = os.getpid()

    def compile_one(src):
        base = os.path.basename(src)[:-3]
        obj = os.path.join(OBJ_DIR, "%s.%d.o" % (base, pid))
        extra = ["-DTSS_SOURCE_DIGEST=\"%s\"" % digest] if base == "api_common" else []
        cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = "%s.%d.tmp" % (LIB_PATH, pid)
    r = subprocess.run([nvcc, "-shared", "-o", tmp] + objs + ["-lcudart"], capture_output=True, text=True)
    for o in objs:
        os.remove(o)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    os.replace(tmp, LIB_PATH)
    with open(STAMP, "w") as fh:
        fh.write(digest)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
