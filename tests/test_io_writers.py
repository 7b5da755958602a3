"""Host-side output writers (lbm_ferrofluid_b200/io.py) against the reference's own functions when /root/reference is
present (build container), and against fixed expectations everywhere."""
import os

import numpy as np
import pytest

REF_IO = "/root/reference/src/LBM/utils/io.py"


def _fields():
    rng = np.random.default_rng(0)
    return (rng.random((2, 1, 3, 4, 5)).astype(np.float32), rng.random((2, 1, 3, 4, 6)).astype(np.float32),
            rng.random((2, 1, 3, 5, 5)).astype(np.float32), rng.random((2, 1, 4, 4, 5)).astype(np.float32))


def test_dump_3d_plt_layout(tmp_path):
    from lbm_ferrofluid_b200.io import dump_3d_plt_file_single

    rho, u, v, w = _fields()
    out = tmp_path / "a.plt"
    dump_3d_plt_file_single(str(out), rho, u, v, w, 1, 0)
    lines = out.read_text().splitlines()
    assert lines[0] == 'TITLE ="Magnetic internal force"' and lines[2] == "ZONE I=5, J=4, K=3" and lines[3] == "F=POINT"
    assert len(lines) == 4 + 3 * 4 * 5
    first = lines[4].split()
    assert first[:3] == ["0", "0", "0"] and float(first[3]) == pytest.approx(float(rho[1, 0, 0, 0, 0]))
    assert float(first[4]) == pytest.approx(0.5 * float(u[1, 0, 0, 0, 0] + u[1, 0, 0, 0, 1]))
    assert lines[5].split()[:3] == ["1", "0", "0"]           # x runs fastest


@pytest.mark.skipif(not os.path.exists(REF_IO), reason="reference sources not present on this machine")
def test_writers_equal_the_reference_byte_for_byte(tmp_path):
    from lbm_ferrofluid_b200.io import dump_2d_plt_file_single, dump_3d_plt_file_single

    src = open(REF_IO).read()
    ns = {}
    exec(src[src.index("def dump_2d_plt_file_single"):], ns)          # the two writers are plain Python (no imports)
    rho, u, v, w = _fields()
    dump_3d_plt_file_single(str(tmp_path / "a.plt"), rho, u, v, w, 1, 0)
    ns["dump_3d_plt_file_single"](str(tmp_path / "b.plt"), rho, u, v, w, 1, 0)
    assert (tmp_path / "a.plt").read_text() == (tmp_path / "b.plt").read_text()
    rng = np.random.default_rng(1)
    d2, uv = rng.random((2, 1, 4, 6)).astype(np.float32), rng.random((2, 2, 4, 6)).astype(np.float32)
    dump_2d_plt_file_single(str(tmp_path / "c.plt"), d2, uv, 1)
    ns["dump_2d_plt_file_single"](str(tmp_path / "d.plt"), d2, uv, 1)
    assert (tmp_path / "c.plt").read_text() == (tmp_path / "d.plt").read_text()
