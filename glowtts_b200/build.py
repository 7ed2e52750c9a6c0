"""Build libglowtts_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m glowtts_b200.build [--force] [--verbose]

The .so lands in glowtts_b200/lib/ (git-ignored, but it travels to the GPU box with the snapshot).
Objects are compiled in parallel, one per .cu, and only when their sources changed.
"""
from __future__ import annotations

import concurrent.futures
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# GTK_LIBDIR / GTK_OBJDIR: build a variant (e.g. GTK_EXTRA_CFLAGS=-DGTK_INSTRUMENT) beside the product library
LIBDIR = os.environ.get("GTK_LIBDIR") or os.path.join(HERE, "lib")
OBJDIR = os.environ.get("GTK_OBJDIR") or os.path.join(HERE, "build")
LIB = os.path.join(LIBDIR, "libglowtts_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
          "-Xptxas", "-v", "-Wno-deprecated-gpu-targets"] + os.environ.get("GTK_EXTRA_CFLAGS", "").split()


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return (glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) +
            glob.glob(os.path.join(CSRC, "*.inl")) + [os.path.join(os.path.dirname(HERE), "include", "glowtts_b200.h")])


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _compile(src, obj, verbose, extra=()):
    cmd = [NVCC, *ARCH, *CFLAGS, *extra, "-c", src, "-o", obj]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    with open(obj + ".log", "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s" % (src, log[-8000:]))
    if verbose:
        print(log)
    return obj


def build(force: bool = False, verbose: bool = False, libdir: str = None, objdir: str = None, extra_cflags=()) -> str:
    libdir, objdir = libdir or LIBDIR, objdir or OBJDIR
    lib = os.path.join(libdir, "libglowtts_b200.so")
    os.makedirs(libdir, exist_ok=True)
    os.makedirs(objdir, exist_ok=True)
    headers = _headers()
    jobs = []
    objs = []
    for src in _sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj))
    if jobs:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for fut in [ex.submit(_compile, s, o, verbose, tuple(extra_cflags)) for s, o in jobs]:
                fut.result()
    if jobs or force or _stale(lib, objs):
        cmd = [NVCC, *ARCH, "-shared", "-o", lib, *objs, "-cudart", "static"]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("link failed:\n" + proc.stdout + proc.stderr)
    return lib


INSTR_DIR = os.path.join(os.path.dirname(HERE), "profiles", "_instr")


def build_instrumented(force: bool = False) -> str:
    """The -DGTK_INSTRUMENT variant (per-CTA globaltimer stamps, cycle counters, gtk_debug_* entry points) beside the
    product library, in profiles/_instr/: loaded through GTK_LIB by the scripts under profiles/ and by bench.py's in-graph
    kernel-span measurement.  Never loaded by the product path."""
    return build(force=force, libdir=INSTR_DIR, objdir=os.path.join(INSTR_DIR, "build"), extra_cflags=("-DGTK_INSTRUMENT",))


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
