"""Builds csrc/*.cu into ldnn_b200/lib/libldnn_b200.so for sm_100a (in-tree; nvcc cross-compiles
without a GPU).  Used by __graft_entry__.build() and by `python -m ldnn_b200.build`."""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(ROOT, "csrc")
LIBDIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIBDIR, "libldnn_b200.so")
OBJDIR = os.path.join(ROOT, "build", "obj")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest() -> str:
    h = hashlib.sha256()
    for d in (CSRC, os.path.join(ROOT, "include")):
        for fn in sorted(os.listdir(d)):
            with open(os.path.join(d, fn), "rb") as f:
                h.update(fn.encode()); h.update(f.read())
    h.update(" ".join(ARCH + FLAGS).encode())
    return h.hexdigest()


def needs_build() -> bool:
    """True when the library is missing or older than csrc/ + include/ (digest stamp).  Without nvcc (a box that only
    received the built .so) an existing library is used as it is."""
    stamp = os.path.join(LIBDIR, "build.stamp")
    if not os.path.exists(LIB):
        return True
    try:
        _nvcc()
    except RuntimeError:
        return False
    try:
        return not (os.path.exists(stamp) and open(stamp).read() == _digest())
    except OSError:
        return False


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(OBJDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, "build.stamp")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    nvcc = _nvcc()
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))

    def compile_one(src):
        obj = os.path.join(OBJDIR, src[:-3] + ".o")
        cmd = [nvcc] + ARCH + FLAGS + ["-I", os.path.join(ROOT, "include"), "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    tmp_lib = f"{LIB}.{os.getpid()}.tmp"          # link beside the target, then rename: readers never see a partial file
    cmd = [nvcc] + ARCH + ["-shared", "-o", tmp_lib] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp_lib, LIB)
    with open(stamp, "w") as f:
        f.write(dig)
    if verbose:
        print(f"[ldnn_b200.build] built {LIB} from {len(srcs)} sources")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
