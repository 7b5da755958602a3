"""CPU checks of the derived marching-cubes case table (lbmff_mc_case_table, host only) and of the numpy restatement
oracle/iso_oracle.py built on it: the properties every correct marching-cubes output has (the reference's PyMCubes is
not available: the triangulation itself is unpinned, see oracle/iso_oracle.py)."""
import numpy as np
import pytest

from oracle import iso_oracle as IO


@pytest.fixture(scope="module")
def table():
    from lbm_ferrofluid_b200.isosurface import case_table
    return case_table()


def test_every_case_uses_exactly_its_cut_edges_in_closed_loops(table):
    ntri, edges = table
    assert ntri[0] == 0 and ntri[255] == 0 and ntri.max() == 5
    for cfg in range(256):
        cut = {e for e in range(12) if ((cfg >> IO.edge_corners(e)[0]) & 1) != ((cfg >> IO.edge_corners(e)[1]) & 1)}
        tris = edges[cfg, : ntri[cfg]]
        assert set(tris.flatten().tolist()) == cut, cfg
        assert (edges[cfg, ntri[cfg]:] == -1).all()
        # directed edges: an interior (fan) edge appears once in each direction, a boundary edge once; the boundary
        # edges form closed loops through every cut edge exactly once and connect cut edges of one cube FACE
        use = {}
        for t in tris:
            for k in range(3):
                use[(t[k], t[(k + 1) % 3])] = use.get((t[k], t[(k + 1) % 3]), 0) + 1
        assert all(v == 1 for v in use.values()), cfg
        boundary = [(a, b) for (a, b) in use if (b, a) not in use]
        assert sorted(a for a, _ in boundary) == sorted(cut) and sorted(b for _, b in boundary) == sorted(cut), cfg
        for a, b in boundary:
            ca, cb = set(IO.edge_corners(a)), set(IO.edge_corners(b))
            corners = ca | cb
            assert any(all(((c >> ax) & 1) == side for c in corners) for ax in range(3) for side in (0, 1)), (cfg, a, b)
        # number of triangles = sum over loops (k - 2)
        assert ntri[cfg] == len(cut) - 2 * _loops(boundary), cfg


def _loops(boundary):
    nxt = dict(boundary)
    seen, n = set(), 0
    for s in nxt:
        if s in seen:
            continue
        n += 1
        e = s
        while e not in seen:
            seen.add(e)
            e = nxt[e]
    return n


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_fields_give_a_watertight_consistently_oriented_surface(table, seed):
    """White noise hits every ambiguous face / cube configuration.  With the volume padded by a constant the surface
    is closed: every directed edge is matched by exactly one opposite edge (2-manifold edges, consistent winding)."""
    ntri, edges = table
    rng = np.random.default_rng(seed)
    u = rng.standard_normal((9, 8, 10)).astype(np.float32)
    u = np.pad(u, 1, constant_values=3.0)
    tri = IO.triangle_soup(u, 0.0, ntri, edges)
    assert len(tri) > 500
    use = IO.edge_use(tri)
    assert all(v == 1 for v in use.values())
    assert all((b, a) in use for (a, b) in use)
    # normals point towards decreasing values, i.e. into the enclosed u < 0 region(s) -> negative signed volume
    assert IO.signed_volume(tri) < 0
    assert IO.signed_volume(IO.triangle_soup(u, 0.0, ntri, edges, flip=True)) > 0


def test_sphere_area_volume_and_vertices_on_the_zero_crossing(table):
    ntri, edges = table
    n, r = 28, 9.3
    z, y, x = np.meshgrid(*(np.arange(n, dtype=np.float64),) * 3, indexing="ij")
    c = (n - 1) / 2 + np.array([0.13, -0.21, 0.34])
    u = (np.sqrt((z - c[0]) ** 2 + (y - c[1]) ** 2 + (x - c[2]) ** 2) - r).astype(np.float32)
    tri = IO.triangle_soup(u, 0.0, ntri, edges, flip=True)      # outward normals for a distance field
    assert abs(IO.signed_volume(tri) / (4 / 3 * np.pi * r ** 3) - 1) < 0.01
    assert abs(IO.area(tri) / (4 * np.pi * r ** 2) - 1) < 0.01
    d = np.linalg.norm(tri.reshape(-1, 3).astype(np.float64) - c, axis=1)
    assert np.abs(d - r).max() < 0.05            # linear interpolation of a distance field along a cube edge
    use = IO.edge_use(tri)
    assert all(v == 1 for v in use.values()) and all((b, a) in use for (a, b) in use)
