"""GPU parity of the tiled soft rasteriser (lbmff_soft_rasterize_*) and of the SoftRenderer host mirror.

Checkers: (1) oracle/raster_oracle.py, the numpy fp32 restatement of the reference forward kernel; (2) when
oracle/_ref/soft_rasterize_ref.so exists (built from the reference's own unmodified .cu by
oracle/build_ref_rasterizer.py; it travels to the GPU box), the reference kernel itself.

Tolerance (fp32, stated here): the image is a sharp sigmoid / softmax of fp32 distances (sigma_val 1e-5, gamma_val
1e-4 by default), the reference is built with --use_fast_math and we are not, so a pixel whose centre lies within
rounding of a triangle edge can land on the other side of a `continue`.  Bar: mean |diff| <= 2e-4 over the image and
at least 99.5 % of the pixels within 2e-3; colours are in [0, 1]."""
import importlib.util
import math
import os

import numpy as np
import pytest
import torch

from oracle import raster_oracle as RO

pytestmark = pytest.mark.gpu

REF_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref", "soft_rasterize_ref.so")
DIST = {"hard": 0, "barycentric": 1, "euclidean": 2}
RGB = {"hard": 0, "softmax": 1}
ALPHA = {"hard": 0, "sum": 1, "prod": 2}


def _close(got, want, what, mean_tol=2e-4, pix_tol=2e-3, frac=0.995):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, what
    assert np.isfinite(got).all(), what
    d = np.abs(got - want)
    assert d.mean() <= mean_tol, (what, "mean", d.mean())
    assert (d <= pix_tol).mean() >= frac, (what, "fraction within tol", (d <= pix_tol).mean())


def _scene(B, F, seed, T=2, duplicate=True):
    rng = np.random.default_rng(seed)
    centre = rng.uniform(-0.9, 0.9, (B, F, 1, 2))
    fv = np.concatenate([centre + rng.uniform(-0.35, 0.35, (B, F, 3, 2)), rng.uniform(1.5, 6.0, (B, F, 3, 1))], axis=3)
    fv[:, : F // 6, :, :2] *= 0.2                         # some tiny faces (sub-pixel: the marching-cubes regime)
    if duplicate:                                        # an exact duplicate with opposite winding (a depth tie: only
        fv[:, -1] = fv[:, -2][:, ::-1]                   # meaningful for the order-independent softmax aggregation)
    tex = rng.uniform(0, 1, (B, F, T, 3))
    return fv.astype(np.float32), tex.astype(np.float32)


def _ref_module():
    if not os.path.exists(REF_SO):
        return None
    spec = importlib.util.spec_from_file_location("soft_rasterize_ref", REF_SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _reference_forward(mod, fv, tex, size, bg, near, far, fill_back, eps, sigma, dist, dist_eps, gamma, rgb, alpha, tt):
    """The call SoftRasterizeFunction.forward makes (renderutils/soft_rasterize.py:60-92)."""
    B, F = fv.shape[:2]
    dev = fv.device
    info = torch.zeros(B, F, 27, device=dev)
    aggrs = torch.zeros(B, 2, size, size, device=dev)
    colors = torch.ones(B, 4, size, size, device=dev)
    for k in range(3):
        colors[:, k] *= bg[k]
    info, aggrs, colors = mod.forward(fv.clone(), tex.clone(), info, aggrs, colors, size, near, far, eps,
                                      sigma, dist, math.log(1.0 / dist_eps - 1.0), gamma, rgb, alpha, tt, fill_back)
    torch.cuda.synchronize()
    return colors, aggrs


CASES = [
    dict(size=48, B=1, F=40, dist="euclidean", rgb="softmax", alpha="prod", tt="surface", sigma=1e-5, gamma=1e-4, fb=True),
    dict(size=40, B=2, F=30, dist="euclidean", rgb="softmax", alpha="prod", tt="surface", sigma=1e-3, gamma=1e-2, fb=False),
    dict(size=33, B=1, F=25, dist="barycentric", rgb="softmax", alpha="sum", tt="vertex", sigma=1e-4, gamma=1e-3, fb=True),
    dict(size=64, B=1, F=50, dist="hard", rgb="hard", alpha="hard", tt="surface", sigma=1e-5, gamma=1e-4, fb=True),
    dict(size=32, B=2, F=20, dist="euclidean", rgb="hard", alpha="prod", tt="vertex", sigma=1e-4, gamma=1e-4, fb=False),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c['dist']}-{c['rgb']}-{c['alpha']}-{c['tt']}-{c['size']}")
def test_rasteriser_against_the_oracle_and_the_reference_kernel(case):
    from lbm_ferrofluid_b200.render import soft_rasterize

    T = 3 if case["tt"] == "vertex" else 2
    fv, tex = _scene(case["B"], case["F"], seed=case["size"], T=T, duplicate=case["rgb"] == "softmax")
    bg = (1.0, 0.5, 0.25)
    got, got_aggr = soft_rasterize(torch.from_numpy(fv).cuda(), torch.from_numpy(tex).cuda(), case["size"], bg, 1.0, 100.0,
                                   case["fb"], 1e-3, case["sigma"], case["dist"], 1e-4, case["gamma"], case["rgb"],
                                   case["alpha"], case["tt"], return_aggrs=True)
    want, want_aggr = RO.soft_rasterize(fv, tex, case["size"], bg, 1.0, 100.0, case["fb"], 1e-3, case["sigma"],
                                        DIST[case["dist"]], 1e-4, case["gamma"], RGB[case["rgb"]], ALPHA[case["alpha"]],
                                        0 if case["tt"] == "surface" else 1)
    _close(got.cpu().numpy(), want, "cuda vs oracle")
    if case["rgb"] == "hard":                             # winning face index per pixel
        assert (got_aggr[:, 1].cpu().numpy() == want_aggr[:, 1]).mean() >= 0.995
    mod = _ref_module()
    if mod is not None:
        ref, _ = _reference_forward(mod, torch.from_numpy(fv).cuda(), torch.from_numpy(tex).cuda(), case["size"], bg, 1.0,
                                    100.0, case["fb"], 1e-3, case["sigma"], DIST[case["dist"]], 1e-4, case["gamma"],
                                    RGB[case["rgb"]], ALPHA[case["alpha"]], 0 if case["tt"] == "surface" else 1)
        _close(want, ref.cpu().numpy(), "oracle vs reference kernel")
        _close(got.cpu().numpy(), ref.cpu().numpy(), "cuda vs reference kernel")


def test_rasteriser_is_deterministic_and_checks_arguments():
    from lbm_ferrofluid_b200.render import soft_rasterize

    fv, tex = _scene(1, 64, seed=7)
    a = soft_rasterize(torch.from_numpy(fv).cuda(), torch.from_numpy(tex).cuda(), 96, (0, 0, 0))
    b = soft_rasterize(torch.from_numpy(fv).cuda(), torch.from_numpy(tex).cuda(), 96, (0, 0, 0))
    assert torch.equal(a, b)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        soft_rasterize(torch.from_numpy(fv), torch.from_numpy(tex).cuda(), 32)
    with pytest.raises(RuntimeError, match=r"\[B, F, 3, 3\]"):
        soft_rasterize(torch.from_numpy(fv).cuda()[:, :, :2], torch.from_numpy(tex).cuda(), 32)
    with pytest.raises(KeyError):
        soft_rasterize(torch.from_numpy(fv).cuda(), torch.from_numpy(tex).cuda(), 32, dist_func="manhattan")
    empty = soft_rasterize(torch.zeros((1, 0, 3, 3), device="cuda"), torch.zeros((1, 0, 2, 3), device="cuda"), 16, (1, 0, 1))
    assert empty[0, 3].abs().max().item() == 0 and empty[0, 0].min().item() == 1 and empty[0, 1].max().item() == 0


def _demo_renderer(size=256):
    from lbm_ferrofluid_b200.render import SoftRenderer

    # demo_3d_LBM_Rosensweig_instability.py:102-113
    eye = torch.tensor([1.6, 0.8, 1.6], device="cuda")
    look_at = torch.tensor([-1.0 / math.sqrt(2.0), 0.0, -1.0 / math.sqrt(2.0)], device="cuda")
    return SoftRenderer(image_size=size, camera_mode="look_at", eye=eye, camera_direction=look_at, near=1.0,
                        bg_color=torch.tensor([1.0, 1.0, 1.0], device="cuda"), device="cuda")


def test_renderer_pipeline_on_a_sphere_mesh():
    """marching cubes -> lighting -> look_at -> perspective -> rasterise, the chain of save_rendered_image
    (io.py:50-84): the silhouette is a disc of the projected radius, the background keeps its colour, a lit sphere
    is brighter on top (light from +y) than below."""
    from lbm_ferrofluid_b200.isosurface import marching_cubes

    n, r = 64, 20.0
    ax = torch.arange(n, device="cuda", dtype=torch.float32)
    z, y, x = torch.meshgrid(ax, ax, ax, indexing="ij")
    u = torch.sqrt((z - 31.5) ** 2 + (y - 31.5) ** 2 + (x - 31.5) ** 2) - r
    verts, faces = marching_cubes(u, 0.0, flip=True)          # a distance field: outward = increasing values
    verts = (verts - n / 2.0) * (2.0 / n)
    tex = 0.68 * torch.ones(1, faces.shape[0], 2, 3, device="cuda")
    ren = _demo_renderer(128)
    img = ren.forward(verts[None], faces[None], tex)
    assert img.shape == (1, 4, 128, 128) and torch.isfinite(img).all()
    alpha = img[0, 3]
    d = math.sqrt(1.6 ** 2 + 0.8 ** 2 + 1.6 ** 2)
    rr = r * 2.0 / n
    r_img = rr / math.sqrt(d * d - rr * rr) / math.tan(math.radians(30.0))      # tangent cone of the sphere, in NDC
    want_area = math.pi * (r_img * 64) ** 2
    assert abs(alpha.sum().item() / want_area - 1) < 0.03
    assert (img[0, :3, 0, 0] == 1).all() and alpha[0, 0] == 0                    # corner = background
    rows = torch.nonzero(alpha.sum(dim=1) > 1)[:, 0]
    top, bottom = rows.min().item(), rows.max().item()
    mid = (top + bottom) // 2
    lit = img[0, 0, top + 3: mid].sum() / alpha[top + 3: mid].sum()
    dark = img[0, 0, mid: bottom - 3].sum() / alpha[mid: bottom - 3].sum()
    assert lit > dark + 0.05


def test_demo_frame_at_256_cubed_matches_the_reference_kernel_and_is_much_faster(tmp_path):
    """The frame of demo_3d_LBM_Rosensweig_instability (256^3 field, 512^2 supersampled raster): our tiled rasteriser
    against the reference's own all-faces-per-pixel kernel on the same projected faces, and save_rendered_image
    end to end (PNG on disk, verts / faces returned)."""
    import torch.nn.functional as F
    from lbm_ferrofluid_b200 import initializers as I
    from lbm_ferrofluid_b200.io import save_rendered_image
    from lbm_ferrofluid_b200.render import soft_rasterize

    res = (256, 256, 256)
    flags = I.shell_flags(res, device="cuda")
    rho = I.rosensweig_rho(res, flags, perturb=6.0)
    phi = -(2.0 * (rho - 0.02381) / (0.2508 - 0.02381) - 1.0)
    ren = _demo_renderer(256)
    out = str(tmp_path / "frame.png")
    verts, faces = save_rendered_image(ren, phi, out, res, 1.0)
    import cv2
    png = cv2.imread(out)
    assert png is not None and png.shape == (256, 256, 3) and png.min() < 200 and png.max() == 255
    assert faces.shape[0] > 100000 and verts.shape[0] == 3 * faces.shape[0]

    # the rasteriser alone, on exactly the faces the renderer projected
    tex = 0.68 * torch.ones(1, faces.shape[0], 2, 3, device="cuda")
    lit = ren.lighting(verts[None], faces[None], tex)
    fv = ren.vertices_to_faces(ren.project_to_image(ren.transform_to_camera_frame(verts[None])), faces[None]).contiguous()
    args = (512, (1.0, 1.0, 1.0), 1.0, 100, True, 1e-3, 1e-5, "euclidean", 1e-4, 1e-4, "softmax", "prod", "surface")
    soft_rasterize(fv, lit, *args)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    ours = soft_rasterize(fv, lit, *args)
    b.record()
    torch.cuda.synchronize()
    ms_ours = a.elapsed_time(b)
    assert torch.isfinite(ours).all() and 0.2 < ours[0, 3].mean().item() < 0.9
    mod = _ref_module()
    if mod is None:
        pytest.skip("oracle/_ref/soft_rasterize_ref.so not built (needs /root/reference at build time)")
    a.record()
    ref, _ = _reference_forward(mod, fv, lit.contiguous(), 512, (1.0, 1.0, 1.0), 1.0, 100.0, True, 1e-3, 1e-5, 2, 1e-4, 1e-4, 1, 2, 0)
    b.record()
    torch.cuda.synchronize()
    ms_ref = a.elapsed_time(b)
    _close(ours.cpu().numpy(), ref.cpu().numpy(), "demo frame: cuda vs reference kernel")
    print(f"\\nsoft rasterise 512^2 x {faces.shape[0]} faces: ours {ms_ours:.2f} ms, reference kernel {ms_ref:.1f} ms "
          f"({ms_ref / ms_ours:.0f}x)")
    assert ms_ours * 10 < ms_ref
