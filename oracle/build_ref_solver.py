"""Test infrastructure: make the REFERENCE's own Python solver and demo scripts available on the GPU box.

    python oracle/build_ref_solver.py

The reference solver (/root/reference/src/LBM/**, pure PyTorch) and its demos (/root/reference/demo/demo_*.py) exist
only in the build container; the GPU box gets a snapshot of this repo.  Same pattern as build_ref_rasterizer.py: the
reference is *compiled from the sources where they lie* -- here with CPython's byte compiler -- into the git-ignored
``oracle/_ref/pyref/`` (sourceless ``.pyc`` files in the legacy layout ``pkg/mod.pyc``, which the import system loads
without the ``.py``), so no reference source enters this repository or its history, yet the unmodified code runs on the
box:

    oracle/_ref/pyref/src/LBM/**.pyc        the solver classes   (import root: oracle/_ref/pyref)
    oracle/_ref/pyref/demo/demo_*.pyc       the demo scripts     (run as ``python demo_x.pyc`` with cwd = a copy of demo/)
    oracle/_ref/pyref/MANIFEST.json         what was compiled, from where, with which interpreter

Consumers (all test / measurement infrastructure, never the product path): oracle/ref_runner.py (the reference run in
a subprocess on CPU or CUDA: parity trajectories and the reference arm of bench.py), tests/test_gpu_reference.py,
tests/test_gpu_demos.py.  The byte code is tied to the interpreter's magic number; build container and GPU box share
one image, and the manifest records the magic so a mismatch is reported instead of an obscure ImportError.
"""
from __future__ import annotations

import glob
import importlib.util
import json
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("LBMFF_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref", "pyref")
MANIFEST = os.path.join(OUT, "MANIFEST.json")


def _sources():
    srcs = sorted(glob.glob(os.path.join(REF, "src", "**", "*.py"), recursive=True))
    demos = sorted(glob.glob(os.path.join(REF, "demo", "demo_*.py")))
    return srcs, demos


def available() -> bool:
    """True when a usable byte-compiled reference is present (built here or shipped with the snapshot)."""
    try:
        with open(MANIFEST) as fh:
            m = json.load(fh)
    except Exception:
        return False
    return m.get("magic") == importlib.util.MAGIC_NUMBER.hex() and os.path.isdir(os.path.join(OUT, "src", "LBM"))


def build(force: bool = False):
    """Byte-compile the reference into oracle/_ref/pyref.  Returns the directory, or None when the reference
    checkout is absent (the GPU box: the prebuilt directory travelled with the snapshot)."""
    if not os.path.isdir(os.path.join(REF, "src", "LBM")):
        return OUT if available() else None
    srcs, demos = _sources()
    newest = max(os.path.getmtime(p) for p in srcs + demos)
    if not force and available() and os.path.getmtime(MANIFEST) >= newest:
        return OUT
    files = []
    for p in srcs + demos:
        rel = os.path.relpath(p, REF)
        dst = os.path.join(OUT, rel[:-3] + ".pyc")
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        # dfile: the path shown in tracebacks -- point at the reference checkout so a failure is attributable
        py_compile.compile(p, cfile=dst, dfile=os.path.join("<reference>", rel), doraise=True,
                           invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
        files.append(rel)
    with open(MANIFEST, "w") as fh:
        json.dump({"reference": REF, "python": sys.version.split()[0], "magic": importlib.util.MAGIC_NUMBER.hex(),
                   "files": files}, fh, indent=1)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
