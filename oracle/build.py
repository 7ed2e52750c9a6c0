"""oracle/build.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Builds the two CPU checkers for the MAS path:

1. ``oracle/_build/libmas_oracle.so`` from ``oracle/mas_oracle.c`` (our plain-C
   restatement of /root/reference/monotonic_align/core.pyx).
2. ``oracle/_ref/core*.so`` -- the reference's OWN ``monotonic_align/core.pyx``
   compiled from the sources where they lie under /root/reference (Cython -> C ->
   gcc, the same steps its setup.py:5-9 performs, without running its build
   system).  Only build OUTPUTS land in ``oracle/_ref/`` (git-ignored); no
   reference source is copied into the repo.  Skipped when /root/reference is
   absent (the GPU box uses the prebuilt files that travelled with the snapshot).

Run:  python oracle/build.py [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")
REF = os.path.join(HERE, "_ref")
REFERENCE_ROOT = "/root/reference"
REF_PYX = os.path.join(REFERENCE_ROOT, "monotonic_align", "core.pyx")


def _run(cmd):
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("command failed: %s\n%s\n%s" % (" ".join(cmd), proc.stdout, proc.stderr))
    return proc


def _stale(target, *sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in sources)


def oracle_lib_path() -> str:
    return os.path.join(BUILD, "libmas_oracle.so")


def ref_core_path() -> str:
    ext = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    return os.path.join(REF, "core" + ext)


def build_oracle(force: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    src = os.path.join(HERE, "mas_oracle.c")
    out = oracle_lib_path()
    if force or _stale(out, src):
        _run(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
              "-o", out, src])
    return out


def build_ref(force: bool = False):
    """Compile the reference's core.pyx in place; returns the .so path or None."""
    out = ref_core_path()
    if not os.path.exists(REF_PYX):
        return out if os.path.exists(out) else None
    os.makedirs(REF, exist_ok=True)
    if not (force or _stale(out, REF_PYX)):
        return out
    import numpy

    c_file = os.path.join(REF, "core.c")
    # Cython refuses to write next to a read-only source; -o sends the C elsewhere.
    _run([sys.executable, "-m", "cython", "-3", REF_PYX, "-o", c_file])
    cflags = (sysconfig.get_config_var("CFLAGS") or "-O2").split()
    inc = [
        "-I" + sysconfig.get_paths()["include"],
        "-I" + numpy.get_include(),
    ]
    _run(["gcc", *cflags, "-fPIC", "-shared", "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
          *inc, "-o", out, c_file])
    # Variant "what prange intended": same source, OpenMP enabled (NOT what ships).
    out_omp = os.path.join(REF, "omp", os.path.basename(out))
    os.makedirs(os.path.dirname(out_omp), exist_ok=True)
    _run(["gcc", *cflags, "-fopenmp", "-fPIC", "-shared",
          "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION", *inc, "-o", out_omp, c_file])
    os.remove(c_file)  # generated from reference source: keep only the binaries
    return out


def main(argv=None):
    force = "--force" in (argv or sys.argv[1:])
    print("oracle:", build_oracle(force))
    print("reference core:", build_ref(force))


if __name__ == "__main__":
    main()
