"""Test infrastructure: make the REFERENCE's own Python solver and demo scripts available on the GPU box.

    python oracle/build_ref_solver.py

The reference solver (/root/reference/src/LBM/**, pure PyTorch) and its demos (/root/reference/demo/demo_*.py) exist
only in the build container; the GPU box gets a snapshot of this repo.  Same pattern as build_ref_rasterizer.py: the
reference is *compiled from the sources where they lie* -- here with CPython's byte compiler -- into ONE git-ignored
archive, ``oracle/_ref/pyref.zip`` (sourceless ``.pyc`` members in the legacy layout ``pkg/mod.pyc``, which zipimport
loads without the ``.py``; an archive because the gpurun snapshot drops loose ``*.pyc`` files), so no reference source
enters this repository or its history, yet the unmodified code runs on the box:

    pyref.zip/src/LBM/**.pyc        the solver classes   (import root: the archive itself, put on sys.path)
    pyref.zip/demo/demo_*.pyc       the demo scripts     (executed by oracle/run_ref_demo.py with cwd = a temp demo/)
    pyref.zip/MANIFEST.json         what was compiled, from where, with which interpreter

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
import tempfile
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("LBMFF_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref", "pyref.zip")


def _sources():
    srcs = sorted(glob.glob(os.path.join(REF, "src", "**", "*.py"), recursive=True))
    demos = sorted(glob.glob(os.path.join(REF, "demo", "demo_*.py")))
    return srcs, demos


def manifest():
    try:
        with zipfile.ZipFile(OUT) as z:
            return json.loads(z.read("MANIFEST.json"))
    except Exception:
        return None


def available() -> bool:
    """True when a usable byte-compiled reference is present (built here or shipped with the snapshot)."""
    m = manifest()
    return bool(m) and m.get("magic") == importlib.util.MAGIC_NUMBER.hex()


def build(force: bool = False):
    """Byte-compile the reference into oracle/_ref/pyref.zip.  Returns the archive path, or None when the reference
    checkout is absent and no archive travelled (the GPU box uses the prebuilt one)."""
    if not os.path.isdir(os.path.join(REF, "src", "LBM")):
        return OUT if available() else None
    srcs, demos = _sources()
    newest = max(os.path.getmtime(p) for p in srcs + demos)
    if not force and available() and os.path.getmtime(OUT) >= newest:
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    files, dirs = [], set()
    with tempfile.TemporaryDirectory() as tmp, zipfile.ZipFile(OUT + ".tmp", "w", zipfile.ZIP_DEFLATED) as z:
        for p in srcs + demos:
            rel = os.path.relpath(p, REF)
            # explicit directory members: zipimport recognises the reference's namespace packages (src/ and src/LBM/
            # have no __init__) only through them
            parts = rel.split(os.sep)[:-1]
            for n in range(1, len(parts) + 1):
                d = "/".join(parts[:n]) + "/"
                if d not in dirs:
                    dirs.add(d)
                    z.writestr(d, "")
            dst = os.path.join(tmp, rel[:-3] + ".pyc")
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            # dfile: the path shown in tracebacks -- point at the reference checkout so a failure is attributable
            py_compile.compile(p, cfile=dst, dfile=os.path.join("<reference>", rel), doraise=True,
                               invalidation_mode=py_compile.PycInvalidationMode.UNCHECKED_HASH)
            z.write(dst, rel[:-3] + ".pyc")
            files.append(rel)
        z.writestr("MANIFEST.json", json.dumps(
            {"reference": REF, "python": sys.version.split()[0], "magic": importlib.util.MAGIC_NUMBER.hex(),
             "files": files}, indent=1))
    os.replace(OUT + ".tmp", OUT)
    return OUT


def demo_code(name: str):
    """Code object of demo/<name>.py from the archive (run_ref_demo.py executes it as __main__)."""
    import marshal
    with zipfile.ZipFile(OUT) as z:
        return marshal.loads(z.read(f"demo/{name}.pyc")[16:])


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
