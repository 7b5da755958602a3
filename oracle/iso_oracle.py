"""Test infrastructure (never imported by the product path): numpy restatement of the device iso-surface extraction
(lbm_ferrofluid_b200/csrc/iso_surface.cu), the CUDA-side replacement of the reference's
``mcubes.marching_cubes`` call (src/LBM/utils/io.py:60-63).  PyMCubes is a third-party dependency that the reference
does not vendor and this image does not have: there is no reference output to pin against -- "parity unpinned" for the
triangulation itself.  What IS checked (tests/test_iso_surface.py) are the properties any correct marching-cubes output
has: every cut edge used, closed oriented loops per cube, a watertight consistently oriented surface on random fields
(all ambiguous cases), vertices on the linear zero crossing, enclosed volume / area of analytic shapes.
"""
import numpy as np


def edge_corners(e):
    """Corner pair (lower index first) of cube edge e = axis*4 + (b0 | b1<<1); corner bit0 = +x, bit1 = +y, bit2 = +z."""
    a, b0, b1 = e // 4, e & 1, (e >> 1) & 1
    o0, o1 = (1 if a == 0 else 0), (1 if a == 2 else 2)
    c0 = (b0 << o0) | (b1 << o1)
    return c0, c0 | (1 << a)


def triangle_soup(u, iso, ntri, edges, flip=False):
    """u [D,H,W] float32 -> [F,3,3] float32 in (axis0, axis1, axis2) coordinates, cell-major order (x fastest)."""
    u = np.asarray(u, dtype=np.float32)
    iso = np.float32(iso)
    D, H, W = u.shape
    out = []
    for z in range(D - 1):
        for y in range(H - 1):
            for x in range(W - 1):
                val = [u[z + (c >> 2), y + ((c >> 1) & 1), x + (c & 1)] for c in range(8)]
                cfg = sum(1 << c for c in range(8) if val[c] < iso)
                for t in range(ntri[cfg]):
                    tri = []
                    for k in range(3):
                        kk = k if flip else (0 if k == 0 else 3 - k)
                        e = int(edges[cfg, t, kk])
                        a = e // 4
                        c0, c1 = edge_corners(e)
                        v0, v1 = val[c0], val[c1]
                        tt = np.float32(np.float32(iso - v0) / np.float32(v1 - v0))
                        tt = np.float32(min(max(tt, np.float32(0)), np.float32(1)))
                        p = [np.float32(x + (c0 & 1)), np.float32(y + ((c0 >> 1) & 1)), np.float32(z + (c0 >> 2))]
                        p[a] = np.float32(p[a] + tt)
                        tri.append([p[2], p[1], p[0]])
                    out.append(tri)
    return np.asarray(out, dtype=np.float32).reshape(-1, 3, 3)


def signed_volume(tri):
    """Divergence theorem: sum of v0 . (v1 x v2) / 6 (positive when the normals point outwards)."""
    t = np.asarray(tri, dtype=np.float64)
    return float(np.einsum("fi,fi->f", t[:, 0], np.cross(t[:, 1], t[:, 2])).sum() / 6.0)


def area(tri):
    t = np.asarray(tri, dtype=np.float64)
    return float(0.5 * np.linalg.norm(np.cross(t[:, 1] - t[:, 0], t[:, 2] - t[:, 0]), axis=1).sum())


def edge_use(tri):
    """{(vertex a, vertex b): count} over DIRECTED triangle edges, vertices keyed by their exact fp32 bits."""
    t = np.ascontiguousarray(np.asarray(tri, dtype=np.float32))
    keys = [[tuple(v.view(np.uint32).tolist()) for v in f] for f in t]
    use = {}
    for f in keys:
        for k in range(3):
            a, b = f[k], f[(k + 1) % 3]
            use[(a, b)] = use.get((a, b), 0) + 1
    return use
