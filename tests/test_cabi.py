"""CPU-side checks of the C-ABI boundary: the library builds, loads, and exports every symbol that
include/lbmff.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import subprocess

import pytest

from lbm_ferrofluid_b200 import _lib, build


def test_header_parses_and_library_exports_every_symbol():
    protos = _lib.parse_header()
    assert len(protos) >= 30
    lib = ctypes.CDLL(build.build())
    for name in protos:
        assert hasattr(lib, name), f"{name} declared in include/lbmff.h but not exported by liblbmff.so"
    assert _lib.load().lbmff_version() >= 100


def test_no_undeclared_exports():
    out = subprocess.run(["nm", "-D", "--defined-only", build.LIB], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T lbmff_" in ln}
    declared = set(_lib.parse_header())
    assert exported == declared, f"undeclared: {exported - declared}; missing: {declared - exported}"


def test_library_is_sm100a_and_torch_free():
    for _ in range(3):     # cuobjdump has been seen to come back empty once on a loaded build machine
        r = subprocess.run(["cuobjdump", "-lelf", build.LIB], capture_output=True, text=True)
        if "sm_100a" in r.stdout:
            break
    assert "sm_100a" in r.stdout, (r.returncode, r.stdout[-300:], r.stderr[-300:])
    ldd = subprocess.run(["ldd", build.LIB], capture_output=True, text=True).stdout
    assert "torch" not in ldd and "c10" not in ldd


def test_bad_arguments_fail_loudly():
    lib = _lib.load()
    rc = lib.lbmff_propagation_d3q19(None, None, 1, 4, 4, 4, None)
    assert rc != 0 and b"lbmff_propagation_d3q19" in lib.lbmff_last_error_string()
    with pytest.raises(RuntimeError):
        _lib.call("lbmff_get_staggered", None, None, 7, 1, 1, 4, 4, None)


def test_ops_refuse_cpu_tensors():
    import torch
    import lbm_ferrofluid_b200 as L

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        L.LBMPropagation2d().propagation(torch.zeros(1, 9, 4, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        L.get_staggered(torch.zeros(1, 2, 4, 4))
