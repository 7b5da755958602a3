"""Test infrastructure (never imported by the product path): numpy restatement of the reference's soft-rasteriser
FORWARD pass -- forward_soft_rasterize_inv_cuda_kernel + forward_soft_rasterize_cuda_kernel
(renderutils/cuda/soft_rasterize_cuda_kernel.cu:236-302 and :305-475, helpers :40-233) as driven by
SoftRasterizeFunction.forward (renderutils/soft_rasterize.py:26-110).  fp32 arithmetic, vectorised over the pixels of
one image, a Python loop over the faces in the reference's order (every pixel visits every face, as the reference's
threads do).  Pinned on the GPU box against the reference kernel itself, compiled unmodified into
oracle/_ref/soft_rasterize_ref.so by oracle/build_ref_rasterizer.py (tests/test_gpu_render.py).
"""
import math

import numpy as np

f32 = np.float32


def face_setup(face):
    """face [3,3] (x, y, z per vertex) -> (inv [3,3], gram [3,3], obtuse [3])  (kernel.cu:236-302)."""
    p = face[:, :2].astype(f32)
    (x0, y0), (x1, y1), (x2, y2) = p
    det = f32(x2 * (y0 - y1) + x0 * (y1 - y2) + x1 * (y2 - y0))
    det = max(det, f32(1e-10)) if det > 0 else min(det, f32(-1e-10))
    adj = np.array([[y1 - y2, x2 - x1, x1 * y2 - x2 * y1],
                    [y2 - y0, x0 - x2, x2 * y0 - x0 * y2],
                    [y0 - y1, x1 - x0, x0 * y1 - x1 * y0]], dtype=f32)
    inv = (adj / det).astype(f32)
    gram = (p @ p.T + f32(1)).astype(f32)
    obt = np.zeros(3, dtype=f32)
    for k in range(3):
        a, b = p[(k + 1) % 3] - p[k], p[(k + 2) % 3] - p[k]
        if a[0] * b[0] + a[1] * b[1] < 0:
            obt[k] = 1
            break
    return inv, gram, obt


def _project(w, gram, a, clamp):
    """Barycentric offset t - w of the closest point on edge (a, a+1) (kernel.cu:92-104 / :137-150)."""
    b, c = (a + 1) % 3, (a + 2) % 3
    d = gram[a] - gram[b]
    t = [None] * 3
    t[a] = (w[0] * d[0] + w[1] * d[1] + w[2] * d[2] - d[b]) / (d[a] - d[b])
    t[b] = f32(1) - t[a]
    t[c] = np.zeros_like(t[a])
    if clamp:
        t = [np.minimum(np.maximum(x, f32(0)), f32(1)) for x in t]
    return [t[k] - w[k] for k in range(3)]


def euclidean(face, gram, obt, w, xp, yp):
    """(sign, squared distance) per pixel (kernel.cu:79-166)."""
    vx, vy = face[:, 0], face[:, 1]
    inside = (w[0] > 0) & (w[1] > 0) & (w[2] > 0) & (w[0] < 1) & (w[1] < 1) & (w[2] < 1)
    best = np.full(xp.shape, f32(100000000.0), dtype=f32)
    for a in range(3):
        t = _project(w, gram, a, clamp=False)
        dx = t[0] * vx[0] + t[1] * vx[1] + t[2] * vx[2]
        dy = t[0] * vy[0] + t[1] * vy[1] + t[2] * vy[2]
        best = np.where(dx * dx + dy * dy < best, dx * dx + dy * dy, best)

    def toward(k, other):  # pixel on the far side of the obtuse corner k: (p - v_k) . (v_other - v_k) > 0
        return (xp - vx[k]) * (vx[other] - vx[k]) + (yp - vy[k]) * (vy[other] - vy[k]) > 0

    n0, n1, n2 = w[0] <= 0, w[1] <= 0, w[2] <= 0
    sel = np.zeros(xp.shape, dtype=np.int64)
    c01 = n1 & n2
    c12 = ~c01 & n2 & n0
    c20 = ~c01 & ~c12 & n0 & n1
    rest = ~c01 & ~c12 & ~c20
    sel = np.where(c01, np.where((obt[0] == 1) & toward(0, 2), 2, 0), sel)
    sel = np.where(c12, np.where((obt[1] == 1) & toward(1, 0), 0, 1), sel)
    sel = np.where(c20, np.where((obt[2] == 1) & toward(2, 1), 1, 2), sel)
    sel = np.where(rest & n0, 1, sel)
    sel = np.where(rest & ~n0 & n1, 2, sel)
    sel = np.where(rest & ~n0 & ~n1 & n2, 0, sel)
    out = np.zeros(xp.shape, dtype=f32)
    for a in range(3):
        t = _project(w, gram, a, clamp=True)
        dx = t[0] * vx[0] + t[1] * vx[1] + t[2] * vx[2]
        dy = t[0] * vy[0] + t[1] * vy[1] + t[2] * vy[2]
        out = np.where(sel == a, dx * dx + dy * dy, out)
    return np.where(inside, f32(1), f32(-1)).astype(f32), np.where(inside, best, out).astype(f32)


def sample(tex, wc, R, mode):
    """[3, ...] colour per pixel (kernel.cu:196-210)."""
    T = tex.shape[0]
    if mode == 0:
        wx, wy = (wc[0] * f32(R)).astype(np.int64), (wc[1] * f32(R)).astype(np.int64)
        lower = (wc[0] + wc[1]) * f32(R) - wx - wy <= 1
        idx = np.where(lower, wy * R + wx, (R - 1 - wy) * R + (R - 1 - wx))
        idx = np.clip(idx, 0, T - 1)
        return np.stack([tex[idx, k] for k in range(3)])
    return np.stack([wc[0] * tex[0, k] + wc[1] * tex[1, k] + wc[2] * tex[2, k] for k in range(3)])


def soft_rasterize(face_vertices, textures, image_size, background=(0, 0, 0), near=1.0, far=100.0, fill_back=True,
                   eps=1e-3, sigma_val=1e-5, dist_func=2, dist_eps=1e-4, gamma_val=1e-4, func_rgb=1, func_alpha=2,
                   texture_type=0):
    """face_vertices [B,F,3,3], textures [B,F,T,3] -> (soft_colors [B,4,is,is], aggrs [B,2,is,is]) in fp32."""
    fv = np.asarray(face_vertices, dtype=f32)
    tx = np.asarray(textures, dtype=f32)
    B, F = fv.shape[:2]
    T = tx.shape[2]
    R = max(int(math.sqrt(T)), 1)
    n = image_size
    near, far, eps, sigma, gamma = f32(near), f32(far), f32(eps), f32(sigma_val), f32(gamma_val)
    threshold = f32(f32(math.log(1.0 / dist_eps - 1.0)) * sigma)
    grow = f32(np.sqrt(threshold))
    rows, cols = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    xp = ((f32(2) * cols.astype(f32) + f32(1) - f32(n)) / f32(n)).astype(f32)
    yp = ((f32(2) * (n - 1 - rows).astype(f32) + f32(1) - f32(n)) / f32(n)).astype(f32)
    colors = np.ones((B, 4, n, n), dtype=f32)
    aggrs = np.zeros((B, 2, n, n), dtype=f32)
    old = np.seterr(all="ignore")
    try:
        for b in range(B):
            ssum = np.full((n, n), np.exp(eps / gamma), dtype=f32)
            smax = np.full((n, n), eps, dtype=f32)
            c = np.stack([np.full((n, n), f32(background[k]), dtype=f32) for k in range(3)])
            if func_rgb == 1:
                c = c * ssum
            alpha = np.full((n, n), f32(1 if func_alpha == 2 else 0), dtype=f32)
            depth_min = np.full((n, n), f32(10000000.0), dtype=f32)
            face_min = np.full((n, n), -1, dtype=np.int64)
            for f in range(F):
                face = fv[b, f]
                inv, gram, obt = face_setup(face)
                act = ~((xp > face[:, 0].max() + grow) | (xp < face[:, 0].min() - grow) |
                        (yp > face[:, 1].max() + grow) | (yp < face[:, 1].min() - grow))
                if not act.any():
                    continue
                w = [(inv[k, 0] * xp + inv[k, 1] * yp + inv[k, 2]).astype(f32) for k in range(3)]
                inside = (w[0] <= 1) & (w[0] >= 0) & (w[1] <= 1) & (w[1] >= 0) & (w[2] <= 1) & (w[2] >= 0)
                if dist_func == 0:
                    frag = inside.astype(f32)
                    act = act & inside
                elif dist_func == 1:
                    d = np.minimum(np.minimum(w[0], w[1]), w[2])
                    d = np.where(d > 0, d * d, -(d * d)).astype(f32)
                    act = act & ~(-d >= threshold)
                    frag = (f32(1) / (f32(1) + np.exp(-d / sigma))).astype(f32)
                else:
                    sign, d = euclidean(face, gram, obt, w, xp, yp)
                    act = act & ~((sign < 0) & (d >= threshold))
                    frag = (f32(1) / (f32(1) + np.exp(-sign * d / sigma))).astype(f32)
                if func_alpha == 0:
                    alpha = np.where(act & (frag > 0.5), f32(1), alpha)
                elif func_alpha == 1:
                    alpha = np.where(act, alpha + frag, alpha)
                else:
                    alpha = np.where(act, alpha * (f32(1) - frag), alpha)
                wc = [np.maximum(np.minimum(x, f32(1)), f32(0)) for x in w]
                wsum = np.maximum(wc[0] + wc[1] + wc[2], f32(1e-5))
                wc = [(x / wsum).astype(f32) for x in wc]
                zp = (f32(1) / (wc[0] / face[0, 2] + wc[1] / face[1, 2] + wc[2] / face[2, 2])).astype(f32)
                act = act & ~((zp < near) | (zp > far))
                front = ((face[2, 1] - face[0, 1]) * (face[1, 0] - face[0, 0])
                         < (face[1, 1] - face[0, 1]) * (face[2, 0] - face[0, 0]))
                col = sample(tx[b, f], wc, R, texture_type)
                if func_rgb == 0:
                    take = act & (zp < depth_min) & inside & (fill_back or front)
                    depth_min = np.where(take, zp, depth_min)
                    face_min = np.where(take, f, face_min)
                    c = np.where(take[None], col, c)
                elif front or fill_back:
                    zn = ((far - zp) / (far - near)).astype(f32)
                    up = act & (zn > smax)
                    rescale = np.where(up, np.exp((smax - zn) / gamma), f32(1)).astype(f32)
                    smax = np.where(up, zn, smax)
                    ez = np.exp((zn - smax) / gamma).astype(f32)
                    ssum = np.where(act, rescale * ssum + ez * frag, ssum).astype(f32)
                    c = np.where(act[None], rescale[None] * c + (ez * frag)[None] * col, c).astype(f32)
            colors[b, 3] = alpha if func_alpha == 0 else (alpha / f32(F) if func_alpha == 1 else f32(1) - alpha)
            if func_rgb == 0:
                bg = np.stack([np.full((n, n), f32(background[k]), dtype=f32) for k in range(3)])
                colors[b, :3] = np.where((face_min != -1)[None], c, bg)
                aggrs[b, 0], aggrs[b, 1] = depth_min, face_min.astype(f32)
            else:
                colors[b, :3] = c / ssum
                aggrs[b, 0], aggrs[b, 1] = ssum, smax
    finally:
        np.seterr(**old)
    return colors, aggrs
