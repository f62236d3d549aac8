"""Radiance RGBE (.hdr) files for the sky image (Sky.IBL_path of render_option.json): the Python mirror of host/hj_hdr.h.
The reference reads the file with stbi_loadf and keeps RGB in a float4 image with w = 0
(/root/reference/include/renderer/texture.h:67-101); stb_image decodes a texel as rgb * 2^(e - 136), no gamma."""
import numpy as np


def float_to_rgbe(rgb):
    """float[H,W,3] -> uint8[H,W,4] (Ward's encoding: shared exponent of the largest component)."""
    rgb = np.asarray(rgb, np.float32)
    m = rgb.max(-1)
    out = np.zeros(rgb.shape[:2] + (4,), np.uint8)
    ok = m > 1e-32
    mant, exp = np.frexp(m[ok])
    scale = (mant * 256.0 / m[ok])[:, None]
    out[ok, :3] = np.clip(rgb[ok] * scale, 0, 255).astype(np.uint8)
    out[ok, 3] = (exp + 128).astype(np.uint8)
    return out


def rgbe_to_float4(rgbe):
    """uint8[H,W,4] -> float32[H,W,4], w = 0: exactly what host/hj_hdr.h (and stb_image) produce."""
    rgbe = np.asarray(rgbe, np.uint8)
    e = rgbe[..., 3].astype(np.int32)
    s = np.where(e > 0, np.ldexp(np.float32(1.0), e - 136), np.float32(0.0)).astype(np.float32)
    out = np.zeros(rgbe.shape[:2] + (4,), np.float32)
    out[..., :3] = rgbe[..., :3].astype(np.float32) * s[..., None]
    return out


def _rle_channel(row):
    out, i, n = bytearray(), 0, len(row)
    while i < n:
        run = 1
        while i + run < n and run < 127 and row[i + run] == row[i]:
            run += 1
        if run >= 4:
            out += bytes([128 + run, row[i]]); i += run
            continue
        j = i
        while j < n and j - i < 128:                      # literal span until the next run of >= 4
            r = 1
            while j + r < n and r < 4 and row[j + r] == row[j]:
                r += 1
            if r >= 4:
                break
            j += 1
        out += bytes([j - i]) + bytes(row[i:j]); i = j
    return bytes(out)


def write_hdr(path, rgb, rle=True):
    """Writes float[H,W,3] as '#?RADIANCE' / FORMAT=32-bit_rle_rgbe / '-Y H +X W', new-style RLE scanlines or flat."""
    rgbe = float_to_rgbe(rgb)
    H, W = rgbe.shape[:2]
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (H, W))
        for y in range(H):
            if rle and 8 <= W < 32768:
                f.write(bytes([2, 2, W >> 8, W & 255]))
                for c in range(4):
                    f.write(_rle_channel(rgbe[y, :, c].tolist()))
            else:
                f.write(rgbe[y].tobytes())
    return rgbe


def procedural_sky(width=64, height=32, seed=4, sun=40.0):
    """A seeded non-constant equirect sky for tests: a vertical gradient, low-frequency colour noise and a small bright 'sun'."""
    rng = np.random.default_rng(seed)
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    grad = (0.2 + 0.8 * (1.0 - v))[:, None, None] * np.array([0.5, 0.7, 1.0], np.float32)
    noise = rng.uniform(0.6, 1.4, (max(2, height // 8), max(2, width // 8), 3)).astype(np.float32)
    noise = np.kron(noise, np.ones((8, 8, 1), np.float32))[:height, :width]
    img = grad * noise
    su, sv = 0.3, 0.22
    d2 = ((u[None, :] - su) * 2.0) ** 2 + (v[:, None] - sv) ** 2
    img = img + sun * np.exp(-d2 / 0.002)[..., None] * np.array([1.0, 0.9, 0.7], np.float32)
    return img.astype(np.float32)
