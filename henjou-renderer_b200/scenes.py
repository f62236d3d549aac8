"""Deterministic procedural stand-ins for the assets BASELINE.json names but the reference snapshot does not ship
(SURVEY §0.4, §8d): the thin-film LUT (`./LUT/Thin_Film_LUT.png`, render_option.json:52), the rtcamp9 scene (C2),
the metamaterial-glass stress scene (C3), the animated sequence (C4) and the ~10 M-triangle scene (C5).
Everything is seeded numpy; geometry is emitted the way the reference's loader emits it (gltfloader.h:1068-1601):
de-indexed triangles (3 vertices each, identity index buffer), one instance per mesh, per-triangle material id."""
import numpy as np

from .scene import SceneData, NodeAnimation, make_material

f32 = np.float32


# ------------------------------------------------------------------------------------------------ thin-film LUT
def _cie_xyz(lam):
    """CIE 1931 colour matching functions, multi-lobe Gaussian fit (Wyman, Sloan, Shirley 2013)."""
    def g(x, mu, s1, s2):
        s = np.where(x < mu, s1, s2)
        return np.exp(-0.5 * ((x - mu) / s) ** 2)
    x = 1.056 * g(lam, 599.8, 37.9, 31.0) + 0.362 * g(lam, 442.0, 16.0, 26.7) - 0.065 * g(lam, 501.1, 20.4, 26.2)
    y = 0.821 * g(lam, 568.8, 46.9, 40.5) + 0.286 * g(lam, 530.9, 16.3, 31.1)
    z = 1.217 * g(lam, 437.0, 11.8, 36.0) + 0.681 * g(lam, 459.0, 26.0, 13.8)
    return x, y, z


def thin_film_lut(size=256, n_film=1.33, n_base=1.5, max_thickness_nm=1000.0):
    """RGBA8 LUT, u = thickness (0..1 -> 0..max nm), v = cos(theta): Airy reflectance of air/film/base integrated
    over 380-780 nm against the CIE observer, converted to linear sRGB (the LUT is bound non-sRGB, renderer.h:854-898)."""
    u = (np.arange(size) + 0.5) / size
    v = (np.arange(size) + 0.5) / size
    d = (u * max_thickness_nm)[None, :, None]                  # [1,U,1]
    c0 = v[:, None, None]                                      # [V,1,1]
    lam = np.linspace(380.0, 780.0, 81)[None, None, :]
    n0, n1, n2 = 1.0, n_film, n_base
    s0 = np.sqrt(np.maximum(0.0, 1 - c0 ** 2))
    c1 = np.sqrt(np.maximum(0.0, 1 - (n0 / n1 * s0) ** 2))
    c2 = np.sqrt(np.maximum(0.0, 1 - (n0 / n2 * s0) ** 2))

    def fres(na, ca, nb, cb):
        rs = (na * ca - nb * cb) / (na * ca + nb * cb)
        rp = (nb * ca - na * cb) / (nb * ca + na * cb)
        return rs, rp
    r01s, r01p = fres(n0, c0, n1, c1)
    r12s, r12p = fres(n1, c1, n2, c2)
    phi = 4 * np.pi * n1 * d * c1 / lam

    def airy(r01, r12):
        num = r01 ** 2 + r12 ** 2 + 2 * r01 * r12 * np.cos(phi)
        den = 1 + (r01 * r12) ** 2 + 2 * r01 * r12 * np.cos(phi)
        return num / den
    R = 0.5 * (airy(r01s, r12s) + airy(r01p, r12p))           # [V,U,L]
    X, Y, Z = _cie_xyz(lam)
    norm = Y.sum()
    xyz = np.stack([(R * X).sum(-1), (R * Y).sum(-1), (R * Z).sum(-1)], -1) / norm
    m = np.array([[3.2406, -1.5372, -0.4986], [-0.9689, 1.8758, 0.0415], [0.0557, -0.2040, 1.0570]])
    rgb = np.clip(xyz @ m.T, 0.0, 1.0)
    # a bare film reflects a few percent; scale so the LUT uses the 8-bit range the way an artist-made F0 LUT does
    rgb = np.clip(rgb * 4.0, 0.0, 1.0)
    out = np.zeros((size, size, 4), np.uint8)
    out[..., :3] = np.round(rgb * 255.0).astype(np.uint8)
    out[..., 3] = 255
    return out


# ------------------------------------------------------------------------------------------------ mesh helpers
def _icosphere_dirs(subdiv):
    """Unit directions of a subdivided icosahedron, de-indexed: float64 [F,3,3]."""
    t = (1 + 5 ** 0.5) / 2
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]])
    tri = v[f]
    for _ in range(subdiv):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = a + b, b + c, c + a
        ab /= np.linalg.norm(ab, axis=1, keepdims=True); bc /= np.linalg.norm(bc, axis=1, keepdims=True); ca /= np.linalg.norm(ca, axis=1, keepdims=True)
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)], 0)
    return tri


def displaced_sphere(subdiv, center, radius, amp=0.0, freq=3.0, phase=(0.0, 0.0, 0.0), stretch=(1.0, 1.0, 1.0)):
    """Sphere with a smooth analytic displacement r(d) = 1 + amp * sum sin(freq d_k + phase_k); normals from the gradient.
    Returns (positions [F*3,3], normals [F*3,3]) float32."""
    d = _icosphere_dirs(subdiv).reshape(-1, 3)
    ph = np.asarray(phase, np.float64)
    s = np.sin(freq * d + ph)
    r = 1.0 + amp * s.sum(1, keepdims=True)
    g = amp * freq * np.cos(freq * d + ph)                       # gradient of the displacement in R^3
    gt = g - (g * d).sum(1, keepdims=True) * d
    n = r * d - gt
    st = np.asarray(stretch, np.float64)
    p = np.asarray(center, np.float64) + radius * r * d * st
    n = n / st
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    return p.astype(f32), n.astype(f32)


def grid_quad(origin, eu, ev, nu, nv):
    """Planar grid of nu x nv cells spanning origin + s*eu + t*ev; normal = normalize(eu x ev)."""
    o, eu, ev = (np.asarray(a, np.float64) for a in (origin, eu, ev))
    s, t = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    s = s.reshape(-1, 1).astype(np.float64); t = t.reshape(-1, 1).astype(np.float64)
    p00 = o + (s / nu) * eu + (t / nv) * ev
    p10 = o + ((s + 1) / nu) * eu + (t / nv) * ev
    p11 = o + ((s + 1) / nu) * eu + ((t + 1) / nv) * ev
    p01 = o + (s / nu) * eu + ((t + 1) / nv) * ev
    tri = np.concatenate([np.stack([p00, p10, p11], 1), np.stack([p00, p11, p01], 1)], 0).reshape(-1, 3)
    n = np.cross(eu, ev); n /= np.linalg.norm(n)
    uv = np.concatenate([np.stack([np.hstack([s / nu, t / nv]), np.hstack([(s + 1) / nu, t / nv]), np.hstack([(s + 1) / nu, (t + 1) / nv])], 1),
                         np.stack([np.hstack([s / nu, t / nv]), np.hstack([(s + 1) / nu, (t + 1) / nv]), np.hstack([s / nu, (t + 1) / nv])], 1)], 0).reshape(-1, 2)
    return tri.astype(f32), np.tile(n.astype(f32), (tri.shape[0], 1)), uv.astype(f32)


def cylinder(base, radius, height, nseg, nring, flute=0.0, nflute=16):
    """Fluted column shaft (open ends), axis +y."""
    a0 = np.arange(nseg) / nseg * 2 * np.pi
    a1 = (np.arange(nseg) + 1) / nseg * 2 * np.pi
    y0 = np.arange(nring) / nring * height
    y1 = (np.arange(nring) + 1) / nring * height
    A0, Y0 = np.meshgrid(a0, y0, indexing="ij"); A1, Y1 = np.meshgrid(a1, y1, indexing="ij")
    A0, Y0, A1, Y1 = (x.reshape(-1) for x in (A0, Y0, A1, Y1))

    def pt(a, y):
        r = radius * (1.0 - flute * (0.5 + 0.5 * np.cos(nflute * a)))
        return np.stack([base[0] + r * np.cos(a), base[1] + y, base[2] + r * np.sin(a)], 1), np.stack([np.cos(a), 0 * a, np.sin(a)], 1)
    p00, n00 = pt(A0, Y0); p10, n10 = pt(A1, Y0); p11, n11 = pt(A1, Y1); p01, n01 = pt(A0, Y1)
    tri = np.concatenate([np.stack([p00, p11, p10], 1), np.stack([p00, p01, p11], 1)], 0).reshape(-1, 3)
    nrm = np.concatenate([np.stack([n00, n11, n10], 1), np.stack([n00, n01, n11], 1)], 0).reshape(-1, 3)
    return tri.astype(f32), nrm.astype(f32)


def box(lo, hi):
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    d = hi - lo
    faces = [(lo, [0, 0, d[2]], [0, d[1], 0]), ([hi[0], lo[1], lo[2]], [0, d[1], 0], [0, 0, d[2]]),
             (lo, [d[0], 0, 0], [0, 0, d[2]]), ([lo[0], hi[1], lo[2]], [0, 0, d[2]], [d[0], 0, 0]),
             (lo, [0, d[1], 0], [d[0], 0, 0]), ([lo[0], lo[1], hi[2]], [d[0], 0, 0], [0, d[1], 0])]
    P, N = [], []
    for o, eu, ev in faces:
        p, n, _ = grid_quad(o, eu, ev, 1, 1)
        P.append(p); N.append(n)
    return np.concatenate(P), np.concatenate(N)


class _Builder:
    def __init__(self):
        self.P, self.N, self.UV, self.M, self.offsets, self.count = [], [], [], [], [], 0
        self.materials = []

    def material(self, **kw):
        self.materials.append(make_material(**kw))
        return len(self.materials) - 1

    def mesh(self, p, n, mat, uv=None):
        """one mesh node = one instance (gltfloader.h:1508-1512); mat: int or per-triangle array"""
        nt = p.shape[0] // 3
        self.offsets.append(self.count)
        self.count += nt
        self.P.append(p); self.N.append(n)
        self.UV.append(uv if uv is not None else np.zeros((p.shape[0], 2), f32))
        self.M.append(np.full(nt, mat, np.uint32) if np.isscalar(mat) else np.asarray(mat, np.uint32))

    def scene(self, sky=(0.8, 0.8, 0.8), lut=None, camera=None):
        s = SceneData()
        s.vertices = np.ascontiguousarray(np.concatenate(self.P), f32)
        s.normals = np.ascontiguousarray(np.concatenate(self.N), f32)
        s.texcoords = np.ascontiguousarray(np.concatenate(self.UV), f32)
        s.indices = np.arange(s.vertices.shape[0], dtype=np.uint32)          # identity (gltfloader.h:1491)
        s.material_ids = np.concatenate(self.M).astype(np.uint32)
        s.prim_offsets = np.asarray(self.offsets, np.uint32)
        s.materials = self.materials
        s.finalize_lights()
        s.set_sky(sky)
        s.lut = lut
        s.animations = [None] * len(self.offsets)
        s.camera = camera
        return s


def _camera(position, target, fov):
    position = np.asarray(position, f32); d = np.asarray(target, np.float64) - position.astype(np.float64)
    d = (d / np.linalg.norm(d)).astype(f32)
    return dict(position=position, direction=d, fov=fov, animation=None)


# ------------------------------------------------------------------------------------------------ C2: rtcamp-like
def rtcamp_like(seed=9, statue_subdiv=8, detail=1.0, n_columns=24, animated=0):
    """Procedural "temple + statue" stand-in for the absent rtcamp9 scene: ~1.5 M triangles at the
    default detail, 30 % thin-film materials (thickness = basecolor.x ~ U[0,1]), emissive ring of 64 quads, constant sky.
    animated > 0: that many orb instances get rigid TRS key tracks (C4)."""
    rng = np.random.default_rng(seed)
    b = _Builder()
    stone = b.material(basecolor=(0.72, 0.70, 0.66), metallic=0.0, roughness=0.6)
    floor_m = b.material(basecolor=(0.35, 0.36, 0.40), metallic=0.0, roughness=0.25)
    gold = b.material(basecolor=(1.0, 0.78, 0.34), metallic=1.0, roughness=0.3)
    light = b.material(basecolor=(1, 1, 1), emission=(24.0, 21.0, 17.0), metallic=0.0, roughness=0.5)
    film = [b.material(basecolor=(float(rng.uniform(0.02, 0.98)), float(rng.uniform(0.2, 0.9)), float(rng.uniform(0.2, 0.9))),
                       metallic=float(rng.choice([0.0, 0.3])), roughness=float(rng.uniform(0.15, 0.5)), is_thinfilm=1) for _ in range(6)]
    plain = [b.material(basecolor=tuple(float(x) for x in rng.uniform(0.2, 0.9, 3)), metallic=0.0, roughness=float(rng.uniform(0.2, 0.9)),
                        sheen=float(rng.choice([0.0, 0.5])), clearcoat=float(rng.choice([0.0, 1.0]))) for _ in range(8)]
    metal = [b.material(basecolor=tuple(float(x) for x in rng.uniform(0.5, 1.0, 3)), metallic=1.0, roughness=float(rng.uniform(0.1, 0.6))) for _ in range(3)]
    glass = b.material(basecolor=(1, 1, 1), metallic=0.0, roughness=0.0, transmission=1.0, ior=1.5)
    # floor, plinth steps, roof slab
    g = max(2, int(160 * detail))
    p, n, uv = grid_quad((-14, 0, -14), (0, 0, 28), (28, 0, 0), g, g); b.mesh(p, n, floor_m, uv)
    for i, (r, h) in enumerate(((4.0, 0.25), (3.2, 0.5), (2.4, 0.75))):
        p, n = box((-r, h - 0.25, -r), (r, h, r)); b.mesh(p, n, stone)
    p, n = box((-11, 9.0, -11), (11, 9.6, 11)); b.mesh(p, n, stone)
    # columns on a ring, fluted
    seg, ring = max(8, int(96 * detail)), max(2, int(48 * detail))
    for k in range(n_columns):
        a = 2 * np.pi * k / n_columns
        cx, cz = 9.5 * np.cos(a), 9.5 * np.sin(a)
        p, n = cylinder((cx, 0.0, cz), 0.55, 9.0, seg, ring, flute=0.12)
        b.mesh(p, n, stone)
    # statue: a stack of displaced blobs, the big one carries most of the triangles; thin-film on 30 % of its faces
    p, n = displaced_sphere(statue_subdiv, (0, 3.2, 0), 1.7, amp=0.09, freq=5.0, phase=(0.3, 1.1, 2.0), stretch=(0.8, 1.35, 0.8))
    nt = p.shape[0] // 3
    cen = p.reshape(-1, 3, 3).mean(1)
    band = (np.floor(cen[:, 1] * 2.5 + 2.0 * np.sin(cen[:, 0] * 3.0)).astype(int)) % 10
    mats = np.where(band < 3, np.asarray(film)[band % len(film)], np.asarray(plain)[band % len(plain)])
    b.mesh(p, n, mats.astype(np.uint32))
    p, n = displaced_sphere(max(statue_subdiv - 2, 1), (0, 6.3, 0), 0.8, amp=0.05, freq=7.0); b.mesh(p, n, gold)
    # orbs between the columns: thin film / metal / glass
    n_orbs = 36
    orb_instances = []
    for k in range(n_orbs):
        a = 2 * np.pi * (k + 0.5) / n_orbs
        rad = float(rng.uniform(4.5, 7.5))
        c = (rad * np.cos(a), float(rng.uniform(0.6, 1.4)), rad * np.sin(a))
        r = float(rng.uniform(0.35, 0.7))
        kind = k % 10
        m = film[k % len(film)] if kind < 3 else (glass if kind == 3 else (metal[k % len(metal)] if kind < 6 else plain[k % len(plain)]))
        sub = max(statue_subdiv - 4, 1)
        if k < animated:      # object space around the origin; the instance transform carries the placement
            p, n = displaced_sphere(sub, (0, 0, 0), r, amp=0.04 * (k % 3), freq=4.0 + k % 5)
            orb_instances.append((len(b.offsets), c))
        else:
            p, n = displaced_sphere(sub, c, r, amp=0.04 * (k % 3), freq=4.0 + k % 5)
        b.mesh(p, n, m)
    # emissive ring of 64 quads under the roof
    for k in range(64):
        a0, a1 = 2 * np.pi * k / 64, 2 * np.pi * (k + 0.7) / 64
        r0, r1 = 6.0, 6.6
        q = np.array([[r0 * np.cos(a0), 8.9, r0 * np.sin(a0)], [r1 * np.cos(a0), 8.9, r1 * np.sin(a0)],
                      [r1 * np.cos(a1), 8.9, r1 * np.sin(a1)], [r0 * np.cos(a1), 8.9, r0 * np.sin(a1)]])
        p = np.stack([q[0], q[1], q[2], q[0], q[2], q[3]]).astype(f32)
        n = np.tile(np.array([[0, -1, 0]], f32), (6, 1))
        b.mesh(p, n, light)
    lut = thin_film_lut(256)
    s = b.scene(sky=(0.8, 0.8, 0.8), lut=lut, camera=_camera((12.5, 3.4, 7.5), (0, 3.0, 0), 0.42))
    for inst, c in orb_instances:
        an = NodeAnimation(t=c)
        for key in range(1, 5):          # keys every 0.25 s: bob + spin
            tt = 0.25 * key
            an.t_key.append(tt); an.t_val.append(np.array([c[0], c[1] + 0.4 * np.sin(2 * np.pi * tt + inst), c[2]], f32))
            ang = np.pi * tt
            an.r_key.append(tt); an.r_val.append(np.array([0, np.sin(ang / 2), 0, np.cos(ang / 2)], f32))
        s.animations[inst] = an
    return s


# ------------------------------------------------------------------------------------------------ C3: metamaterial glass
def metaglass_room(seed=3, subdiv=6):
    """Closed room + nested negative-IOR spheres and slabs (ior 1.5, ideal_specular): deep, divergent paths."""
    rng = np.random.default_rng(seed)
    b = _Builder()
    white = b.material(basecolor=(0.8, 0.8, 0.8), metallic=0.0, roughness=0.7)
    red = b.material(basecolor=(0.8, 0.1, 0.1), metallic=0.0, roughness=0.7)
    green = b.material(basecolor=(0.1, 0.8, 0.1), metallic=0.0, roughness=0.7)
    light = b.material(basecolor=(1, 1, 1), emission=(18, 18, 18), metallic=0.0)
    glass = b.material(basecolor=(1, 1, 1), metallic=0.0, roughness=0.0, transmission=1.0, ior=1.5)
    checker = b.material(basecolor=(0.3, 0.5, 0.9), metallic=0.0, roughness=0.4)
    for (o, eu, ev, m) in (((-4, 0, -4), (0, 0, 8), (8, 0, 0), white), ((-4, 6, -4), (8, 0, 0), (0, 0, 8), white),
                           ((-4, 0, -4), (8, 0, 0), (0, 6, 0), white), ((-4, 0, 4), (0, 6, 0), (8, 0, 0), checker),
                           ((-4, 0, -4), (0, 6, 0), (0, 0, 8), red), ((4, 0, -4), (0, 0, 8), (0, 6, 0), green)):
        p, n, uv = grid_quad(o, eu, ev, 8, 8); b.mesh(p, n, m, uv)
    p, n, uv = grid_quad((-1.2, 5.98, -1.2), (2.4, 0, 0), (0, 0, 2.4), 2, 2); b.mesh(p, n, light, uv)
    for r in (1.6, 1.1, 0.6):                                   # nested shells
        p, n = displaced_sphere(subdiv, (0, 2.4, 0), r); b.mesh(p, n, glass)
    for k in range(3):                                          # slabs
        x = -2.8 + 0.5 * k
        p, n = box((x, 0.2, -2.0 + 0.3 * k), (x + 0.15, 4.0, 1.5 + 0.3 * k)); b.mesh(p, n, glass)
    for k in range(6):
        c = (float(rng.uniform(1.5, 3.2)), float(rng.uniform(0.5, 4.5)), float(rng.uniform(-3, 3)))
        p, n = displaced_sphere(max(subdiv - 2, 1), c, float(rng.uniform(0.3, 0.6))); b.mesh(p, n, glass if k % 2 else checker)
    return b.scene(sky=(0.0, 0.0, 0.0), camera=_camera((0.0, 3.0, -3.8), (0, 2.4, 0), 0.6))


# ------------------------------------------------------------------------------------------------ C5: many triangles
def icosphere_field(n_spheres=2000, subdiv=4, seed=5):
    """~10 M triangles at the defaults (2000 x 5120): displaced icospheres in a box with an emissive ceiling patch."""
    rng = np.random.default_rng(seed)
    b = _Builder()
    mats = [b.material(basecolor=tuple(float(x) for x in rng.uniform(0.2, 0.95, 3)), metallic=float(m), roughness=float(rng.uniform(0.15, 0.8)),
                       is_thinfilm=int(t)) for m, t in ((0, 0), (0, 0), (0, 1), (1, 0), (0, 0), (0, 1), (1, 0), (0, 0))]
    wall = b.material(basecolor=(0.7, 0.7, 0.7), metallic=0.0, roughness=0.8)
    light = b.material(basecolor=(1, 1, 1), emission=(30, 30, 30), metallic=0.0)
    L = 20.0
    for (o, eu, ev) in (((-L, 0, -L), (0, 0, 2 * L), (2 * L, 0, 0)), ((-L, 0, -L), (2 * L, 0, 0), (0, L, 0)), ((-L, 0, -L), (0, L, 0), (0, 0, 2 * L)),
                        ((L, 0, -L), (0, 0, 2 * L), (0, L, 0))):
        p, n, uv = grid_quad(o, eu, ev, 16, 16); b.mesh(p, n, wall, uv)
    p, n, uv = grid_quad((-6, L - 0.01, -6), (12, 0, 0), (0, 0, 12), 4, 4); b.mesh(p, n, light, uv)
    base = _icosphere_dirs(subdiv).reshape(-1, 3)
    for k in range(n_spheres):
        c = rng.uniform((-L + 1.5, 0.8, -L + 1.5), (L - 1.5, L - 3.0, L - 1.5))
        r = rng.uniform(0.35, 0.9)
        amp, freq, ph = rng.uniform(0.0, 0.08), rng.uniform(2, 8), rng.uniform(0, 6.28, 3)
        s = np.sin(freq * base + ph)
        rr = 1.0 + amp * s.sum(1, keepdims=True)
        g = amp * freq * np.cos(freq * base + ph)
        gt = g - (g * base).sum(1, keepdims=True) * base
        n = rr * base - gt
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        b.mesh((c + r * rr * base).astype(f32), n.astype(f32), mats[k % len(mats)])
    lut = thin_film_lut(256)
    return b.scene(sky=(0.5, 0.6, 0.8), lut=lut, camera=_camera((-L + 1.0, L * 0.55, -L + 1.0), (0, L * 0.3, 0), 0.55))


# ------------------------------------------------------------------------------------------------ furnace
def furnace_sphere(metallic=1.0, roughness=0.5, subdiv=5, basecolor=(1.0, 1.0, 1.0), sky=1.0):
    """White-furnace test of the reference's default config (render_option.json:11-12, asset missing): one sphere
    under a constant sky."""
    b = _Builder()
    m = b.material(basecolor=basecolor, metallic=metallic, roughness=roughness)
    p, n = displaced_sphere(subdiv, (0, 0, 0), 1.0); b.mesh(p, n, m)
    return b.scene(sky=(sky, sky, sky), camera=_camera((0, 0, 4.0), (0, 0, 0), 0.5))
