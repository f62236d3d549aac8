"""TEST INFRASTRUCTURE — CPU restatement (numpy, float32, operation for operation) of the post stage
henjou-renderer_b200/csrc/hj_post.cu. Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.

What it stands for in the reference: OptixDenoiserManager::denoise (/root/reference/include/renderer/denoiser.h:143-177),
driven per frame by renderer.h:1256-1283, and float4ConvertColor (renderer.h:73-101). The OptiX AI denoiser is a closed
network with no source: PARITY UNPINNED for the two denoising modes — this file pins the stand-in (edge-avoiding a-trous
filter, Dammertz et al. 2010, guided by the albedo and normal AOVs) against itself on the CPU; the pass-through mode
(Render_mode "Default", blendFactor 1) is exact against the colour AOV.
"""
import numpy as np

F = np.float32
H_TAP = {0: F(0.375), 1: F(0.25), 2: F(0.0625)}


def _dist2(a, b):
    d = a[..., :3] - b[..., :3]
    return (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]


def atrous_pass(color, albedo, normal, step, inv_sc2, inv_sn2, inv_sa2):
    """One pass of k_atrous: 25 taps with holes `step`, taps outside the frame skipped, dy outer / dx inner."""
    Hh, Ww = color.shape[:2]
    s = np.zeros((Hh, Ww, 3), F)
    ws = np.zeros((Hh, Ww), F)
    with np.errstate(invalid="ignore"):
        ok = np.isfinite((color[..., 0] + color[..., 1]) + color[..., 2])       # non-finite taps are skipped; a non-finite centre
    safe = np.where(ok[..., None], color, F(0)).astype(F)                      # is rebuilt from the guides alone
    lum = np.maximum((F(0.2126) * safe[..., 0] + F(0.7152) * safe[..., 1]) + F(0.0722) * safe[..., 2], F(0))
    inv_sc2_p = (inv_sc2 / ((F(1) + lum) * (F(1) + lum))).astype(F)            # colour differences relative to 1 + centre luminance
    for dy in range(-2, 3):
        for dx in range(-2, 3):
            oy, ox = dy * step, dx * step
            y0, y1 = max(0, -oy), min(Hh, Hh - oy)
            x0, x1 = max(0, -ox), min(Ww, Ww - ox)
            if y0 >= y1 or x0 >= x1:
                continue
            P = (slice(y0, y1), slice(x0, x1))
            Q = (slice(y0 + oy, y1 + oy), slice(x0 + ox, x1 + ox))
            dc = np.where(ok[P], _dist2(safe[P], safe[Q]), F(0)).astype(F)
            e = (dc * inv_sc2_p[P] + _dist2(normal[P], normal[Q]) * inv_sn2) + _dist2(albedo[P], albedo[Q]) * inv_sa2
            w = np.where(ok[Q], np.exp(-e).astype(F) * (H_TAP[abs(dx)] * H_TAP[abs(dy)]), F(0)).astype(F)
            for k in range(3):
                s[P + (k,)] += safe[Q + (k,)] * w
            ws[P] += w
    out = color.copy()
    has = ws > 0
    out[has, :3] = s[has] / ws[has][:, None]
    return out


def denoise(color, albedo, normal, iterations=5, sigma_color=2.0, sigma_normal=0.1, sigma_albedo=0.05):
    color, albedo, normal = (np.ascontiguousarray(a, F) for a in (color, albedo, normal))
    sc, sn, sa = F(1e-3) * F(round(sigma_color * 1000)), F(1e-3) * F(round(sigma_normal * 1000)), F(1e-3) * F(round(sigma_albedo * 1000))
    for i in range(iterations):
        sci = sc / F(1 << i)
        color = atrous_pass(color, albedo, normal, 1 << i, F(1) / (sci * sci), F(1) / (sn * sn), F(1) / (sa * sa))
    return color


def upscale2x(img):
    """k_upscale2x: bilinear, pixel centres aligned, clamped."""
    img = np.ascontiguousarray(img, F)
    Hh, Ww = img.shape[:2]

    def axis(n):
        f = (np.arange(2 * n, dtype=F) + F(0.5)) * F(0.5) - F(0.5)
        fl = np.floor(f)
        return np.maximum(fl.astype(np.int64), 0), np.minimum(fl.astype(np.int64) + 1, n - 1), (f - fl).astype(F)
    y0, y1, ty = axis(Hh)
    x0, x1, tx = axis(Ww)
    tx = tx[None, :, None]; ty = ty[:, None, None]
    top = img[y0][:, x0] + (img[y0][:, x1] - img[y0][:, x0]) * tx
    bot = img[y1][:, x0] + (img[y1][:, x1] - img[y1][:, x0]) * tx
    return (top + (bot - top) * ty).astype(F)


def to_rgba8(img):
    """float4ConvertColor (renderer.h:73-101): toSRGB (threshold 0.0031308, exponent 1/2.4), min(255, uint(x*256))."""
    c = np.ascontiguousarray(img[..., :3], F)
    lo = c * F(12.92)
    hi = F(1.055) * np.power(np.maximum(c, F(0)), F(1) / F(2.4), dtype=F) - F(0.055)
    s = np.where(c < F(0.0031308), lo, hi).astype(F)
    q = np.minimum(255, np.maximum(s * F(256.0), F(0)).astype(np.uint32)).astype(np.uint8)
    return np.concatenate([q, np.full(q.shape[:2] + (1,), 255, np.uint8)], -1)
