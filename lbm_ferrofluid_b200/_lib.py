"""ctypes binding of liblbmff.so, generated from include/lbmff.h (the single source of truth).

There is NO CPU fallback: if the library cannot be loaded, every op raises.
"""
from __future__ import annotations

import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(HERE, "..", "include", "lbmff.h")
LIB_PATH = os.path.join(HERE, "liblbmff.so")


class FluidParams(ctypes.Structure):
    """Mirror of ``lbmff_fluid_params`` (include/lbmff.h)."""
    _fields_ = [(n, ctypes.c_double) for n in (
        "dx", "dt", "density_liquid", "density_gas", "rho_liquid", "rho_gas", "kappa", "tau", "tau_f", "tau_g",
        "gravity", "cot_theta")]


class Slab(ctypes.Structure):
    """Mirror of ``lbmff_slab`` (include/lbmff.h, slab tier)."""
    _fields_ = [(n, ctypes.c_int) for n in ("Dg", "zoff", "zb", "zn")]


_TYPE_MAP = [
    (re.compile(r"^const\s+lbmff_\w+_params\s*\*$"), None),  # filled per struct below
    (re.compile(r"^(const\s+)?(float|double|uint8_t|void|int)\s*\*$"), ctypes.c_void_p),
    (re.compile(r"^lbmff_stream_t$"), ctypes.c_void_p),
    (re.compile(r"^int$"), ctypes.c_int),
    (re.compile(r"^long long$"), ctypes.c_longlong),
    (re.compile(r"^double$"), ctypes.c_double),
    (re.compile(r"^float$"), ctypes.c_float),
]

_STRUCTS = {"lbmff_fluid_params": FluidParams, "lbmff_slab": Slab}


def register_struct(name, cls):
    _STRUCTS[name] = cls


def parse_header(path: str = HEADER):
    """Return {function name: (restype, [argtypes])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(long long|int|const char\s*\*)\s+(lbmff_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"^(.*?)(\w+)$", a)           # strip the parameter name
                ty = mm.group(1).strip() if mm else a
                sm = re.match(r"^const\s+(lbmff_\w+)\s*\*$", ty)
                if sm and sm.group(1) in _STRUCTS:
                    argtypes.append(ctypes.POINTER(_STRUCTS[sm.group(1)]))
                    continue
                for rx, ct in _TYPE_MAP[1:]:
                    if rx.match(ty):
                        argtypes.append(ct)
                        break
                else:
                    raise RuntimeError(f"lbmff.h: cannot map C type '{ty}' of {name}")
        protos[name] = (ctypes.c_char_p if "char" in ret else (ctypes.c_longlong if "long" in ret else ctypes.c_int),
                        argtypes)
    return protos


_lib = None


def load():
    """Load liblbmff.so (rebuilding it with nvcc first when it is missing or older than csrc/ / include/)."""
    global _lib
    if _lib is not None:
        return _lib
    from . import build as _build
    if _build.needs_build():
        # missing, or older than a source / header: never run a stale binary against an edited csrc/ (the GPU box
        # has nvcc too).  Without nvcc a missing library is fatal, a stale one only warned about.
        if _build.find_nvcc() is not None:
            _build.build()
        elif not os.path.exists(LIB_PATH):
            raise RuntimeError("liblbmff.so is missing and nvcc was not found: build it with "
                               "`python -m lbm_ferrofluid_b200.build` (there is no CPU fallback)")
        else:
            import warnings
            warnings.warn("liblbmff.so is older than its sources and nvcc is not available to rebuild it")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in parse_header().items():
        fn = getattr(lib, name)   # AttributeError here = header/library mismatch: fail loudly
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def call(name: str, *args):
    """Call an int-returning entry point; raise RuntimeError with the library's message on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.lbmff_last_error_string()
        raise RuntimeError(f"{name} failed (code {rc}): {msg.decode() if msg else '?'}")
