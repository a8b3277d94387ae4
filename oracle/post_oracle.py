"""TEST INFRASTRUCTURE ONLY — CPU restatement (numpy, float32) of the film post-processing in
final-stage-path-tracer-2-0_b200/csrc/fs2_post.cuh.

Parity unpinned, by construction: the reference describes denoise / bloom / flare in its README
but has no code for them, so there is nothing upstream to pin against.  The
definition is this repo's; the guides it reads are the reference's DisplayFrame planes
(source/frame.h:86-91).  Only tests/ may import this module.

All images are (h, w, ...) arrays, row 0 = bottom scanline (the film's order).
"""
import numpy as np

F = np.float32
TAP = np.array([0.375, 0.25, 0.0625], F)


def lum(c):
    return (F(0.2126) * c[..., 0] + F(0.7152) * c[..., 1]) + F(0.0722) * c[..., 2]


def _shift(a, dy, dx, fill=0):
    """b[y, x] = a[y + dy, x + dx] where that is inside, else fill; second result = inside."""
    h, w = a.shape[:2]
    b = np.full_like(a, fill)
    ok = np.zeros((h, w), bool)
    y0, y1 = max(0, -dy), min(h, h - dy)
    x0, x1 = max(0, -dx), min(w, w - dx)
    if y0 < y1 and x0 < x1:
        b[y0:y1, x0:x1] = a[y0 + dy:y1 + dy, x0 + dx:x1 + dx]
        ok[y0:y1, x0:x1] = True
    return b, ok


def sigma_plane(l, ids):
    s1 = np.zeros_like(l)
    s2 = np.zeros_like(l)
    cnt = np.zeros_like(l)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            lq, ok = _shift(l, dy, dx)
            iq, _ = _shift(ids, dy, dx)
            m = ok & (iq == ids)
            s1 = np.where(m, s1 + lq, s1)
            s2 = np.where(m, s2 + lq * lq, s2)
            cnt = np.where(m, cnt + F(1), cnt)
    m1, m2 = s1 / cnt, s2 / cnt
    return np.sqrt(np.maximum(m2 - m1 * m1, F(0))).astype(F)


def atrous(col, normal, depth, ids, sigma, step, scale, p):
    l = lum(col)
    has_n = (normal[..., 0] * normal[..., 0] + normal[..., 1] * normal[..., 1]
             + normal[..., 2] * normal[..., 2]) > 0
    inv_z = F(1) / (F(p.sigma_depth) * np.maximum(np.abs(depth), F(1e-3)))
    inv_l = F(1) / (F(p.sigma_colour) * sigma * F(scale) + F(1e-4))
    acc = np.zeros_like(col)
    aw = np.zeros_like(l)
    for dy in range(-2, 3):
        for dx in range(-2, 3):
            cq, ok = _shift(col, dy * step, dx * step)
            nq, _ = _shift(normal, dy * step, dx * step)
            zq, _ = _shift(depth, dy * step, dx * step)
            iq, _ = _shift(ids, dy * step, dx * step)
            lq, _ = _shift(l, dy * step, dx * step)
            wt = np.full_like(l, TAP[abs(dx)] * TAP[abs(dy)])
            d = np.maximum((normal[..., 0] * nq[..., 0] + normal[..., 1] * nq[..., 1])
                           + normal[..., 2] * nq[..., 2], F(0))
            wn = np.power(d, F(p.sigma_normal), dtype=F)
            wt = np.where(has_n, wt * wn, wt)
            wt = wt * np.exp(-np.abs(depth - zq) * inv_z, dtype=F)
            wt = wt * np.exp(-np.abs(l - lq) * inv_l, dtype=F)
            wt = np.where(ok & (iq == ids), wt, F(0)).astype(F)
            acc = acc + wt[..., None] * cq
            aw = aw + wt
    return (acc / aw[..., None]).astype(F)


def bright(col, threshold):
    l = lum(col)
    k = np.maximum(l - F(threshold), F(0)) / np.maximum(l, F(1e-6))
    return (col * k[..., None]).astype(F)


def blur(img, radius, axis):
    sg = F(radius) / F(3)
    inv = F(1) / (F(2) * sg * sg)
    n = img.shape[axis]
    acc = np.zeros_like(img)
    aw = F(0)
    idx = np.arange(n)
    for k in range(-radius, radius + 1):
        wt = np.exp(-F(k * k) * inv, dtype=F)
        acc = acc + wt * np.take(img, np.clip(idx + k, 0, n - 1), axis=axis)
        aw = F(aw + wt)
    return (acc / aw).astype(F)


def post(mean, normal, depth, ids, p):
    """mean (h,w,3), normal (h,w,3), depth (h,w), ids (h,w); p = abi.fs_post_params.
    Returns (rgb float32 (h,w,3), rgb8 uint8 (h,w,3))."""
    col = mean.astype(F)
    h, w = depth.shape
    if p.denoise_passes:
        sigma = sigma_plane(lum(col), ids)
        scale = 1.0
        for i in range(p.denoise_passes):
            col = atrous(col, normal, depth, ids, sigma, 1 << i, scale, p)
            scale *= 0.5
    out = col.copy()
    glow_on = p.bloom_strength > 0 or (p.flare_strength > 0 and p.flare_ghosts > 0)
    if glow_on:
        g = blur(blur(bright(col, p.bloom_threshold), p.bloom_radius, 1), p.bloom_radius, 0)
        if p.bloom_strength > 0:
            out = out + F(p.bloom_strength) * g
        if p.flare_strength > 0 and p.flare_ghosts > 0:
            cx, cy = F(0.5) * F(w - 1), F(0.5) * F(h - 1)
            k = F(p.flare_strength) / F(p.flare_ghosts)
            xs = np.arange(w, dtype=F)[None, :].repeat(h, 0)
            ys = np.arange(h, dtype=F)[:, None].repeat(w, 1)
            for i in range(1, p.flare_ghosts + 1):
                t = F(i) / F(p.flare_ghosts)
                qx = np.floor((cx + (cx - xs) * t) + F(0.5)).astype(np.int64)
                qy = np.floor((cy + (cy - ys) * t) + F(0.5)).astype(np.int64)
                ok = (qx >= 0) & (qy >= 0) & (qx < w) & (qy < h)
                e = g[np.clip(qy, 0, h - 1), np.clip(qx, 0, w - 1)]
                out = out + np.where(ok[..., None], k * e, F(0))
    out = out.astype(F)
    sat = np.clip(out.astype(np.float64), 0.0, 1.0)
    rgb8 = (255.0 * np.power(sat, 1.0 / 2.2) + 0.5).astype(np.uint8)
    return out, rgb8
