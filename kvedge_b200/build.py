"""In-tree build of libkvedge_b200.so (nvcc, sm_100a only).

``python -m kvedge_b200.build [--tools] [--force] [-v]`` or ``__graft_entry__.build()``.

Two libraries come out of the same sources:

* ``libkvedge_b200.so`` -- the product.  It contains the product kernels only and reads no
  environment variable.
* ``libkvedge_b200_tools.so`` (``-DKVEDGE_TOOLS``) -- additionally holds the sweep shapes, the
  traffic-only skeleton kernel and the ``KVEDGE_*`` tuning knobs that ``tools/`` uses.  Nothing
  under ``kvedge_b200/`` or ``tests/`` loads it unless ``_lib.use_tools_build()`` is called.

Both are built next to this file so that they travel to the GPU box with the repo snapshot;
they are git-ignored.  Objects are cached per source under ``kvedge_b200/build/``.
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libkvedge_b200.so")
LIB_TOOLS = os.path.join(HERE, "libkvedge_b200_tools.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall,-pthread",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _headers():
    return sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.hpp")) +
                  glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; libkvedge_b200.so cannot be built")
    return nvcc


def _run(cmd, verbose):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
        raise RuntimeError("nvcc failed: " + os.path.basename(cmd[-1]))
    if verbose:
        sys.stderr.write(res.stderr)


def _object_for(src: str, tools: bool, force: bool, verbose: bool) -> str:
    """Compile one source unless an object built from the same bytes (source, headers, flags)
    is already there."""
    h = hashlib.sha256()
    for dep in [src] + _headers() + [os.path.abspath(__file__)]:
        with open(dep, "rb") as f:
            h.update(f.read())
    h.update(repr((NVCC_FLAGS, tools)).encode())
    kind = "tools" if tools else "product"
    obj = os.path.join(OBJ, "%s.%s.%s.o" % (os.path.basename(src), kind, h.hexdigest()[:16]))
    if force or not os.path.exists(obj):
        os.makedirs(OBJ, exist_ok=True)
        for stale in glob.glob(os.path.join(OBJ, "%s.%s.*.o" % (os.path.basename(src), kind))):
            os.remove(stale)
        cmd = [_nvcc()] + NVCC_FLAGS + (["-DKVEDGE_TOOLS"] if tools else []) + \
            (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        _run(cmd, verbose)
    return obj


def build_native(force: bool = False, verbose: bool = False, tools: bool = False) -> str:
    """Compile every CUDA/C++ source under csrc/ into one shared library."""
    lib = LIB_TOOLS if tools else LIB
    if not force and os.path.exists(lib):
        # fast path (also what the GPU box takes: the library travels there, the object cache
        # does not): nothing is newer than the library
        t = os.path.getmtime(lib)
        if all(os.path.getmtime(d) <= t for d in sources() + _headers() + [os.path.abspath(__file__)]):
            return lib
    with ThreadPoolExecutor(max(1, min(len(sources()), os.cpu_count() or 1))) as ex:
        objs = list(ex.map(lambda s: _object_for(s, tools, force, verbose), sources()))
    _run([_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-pthread",
          "-o", lib + ".tmp"] + objs, verbose)
    os.replace(lib + ".tmp", lib)
    return lib


if __name__ == "__main__":
    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv, tools="--tools" in sys.argv))
