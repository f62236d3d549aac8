"""Generates the polynomial coefficients of henjou-renderer_b200/csrc/hj_libm.h: interpolation at Chebyshev nodes
(near-minimax) of the reduced-range kernels, evaluated with mpmath at 50 digits and rounded to double.
    python tools/libm_coeffs.py"""
import mpmath as mp

mp.mp.dps = 50


def cheb_fit(fn, a, b, deg):
    n = deg + 1
    xs = [(a + b) / 2 + (b - a) / 2 * mp.cos(mp.pi * (2 * k + 1) / (2 * n)) for k in range(n)]
    A = mp.matrix(n, n)
    y = mp.matrix(n, 1)
    for i, x in enumerate(xs):
        for j in range(n):
            A[i, j] = x ** j
        y[i] = fn(x)
    c = mp.lu_solve(A, y)
    coef = [float(c[j]) for j in range(n)]
    # max error of the double-rounded polynomial on a dense grid
    err = 0
    for k in range(2001):
        x = a + (b - a) * mp.mpf(k) / 2000
        p = sum(mp.mpf(coef[j]) * x ** j for j in range(n))
        err = max(err, abs(p - fn(x)))
    return coef, float(err)


def show(name, coef, err):
    print("// %s: max abs error of the kernel %.2e" % (name, err))
    print("   ", ", ".join(repr(c) for c in coef))


def lim(f, x, x0):
    return f(x) if x != 0 else x0


# sin(r) = r + r^3 S(z), z = r^2 in [0, (pi/4)^2]
zmax = (mp.pi / 4 + mp.mpf("0.001")) ** 2
show("S", *cheb_fit(lambda z: (mp.sin(mp.sqrt(z)) / mp.sqrt(z) - 1) / z if z else mp.mpf(-1) / 6, 0, zmax, 5))
# cos(r) = 1 - z/2 + z^2 C(z)
show("C", *cheb_fit(lambda z: (mp.cos(mp.sqrt(z)) - 1 + z / 2) / z ** 2 if z else mp.mpf(1) / 24, 0, zmax, 5))
# asin(x) = x + x z R(z), z = x^2 in [0, 0.25]
show("A", *cheb_fit(lambda z: (mp.asin(mp.sqrt(z)) / mp.sqrt(z) - 1) / z if z else mp.mpf(1) / 6, 0, mp.mpf("0.2501"), 11))
# atan(u) = u + u z T(z), z = u^2 in [0, (sqrt2-1)^2]
show("T", *cheb_fit(lambda z: (mp.atan(mp.sqrt(z)) / mp.sqrt(z) - 1) / z if z else mp.mpf(-1) / 3, 0, (mp.sqrt(2) - 1) ** 2 + mp.mpf("1e-4"), 10))
# log(m) = 2 s (1 + z L(z)), s = (m-1)/(m+1), z = s^2 in [0, 0.02944]
show("L", *cheb_fit(lambda z: (mp.atanh(mp.sqrt(z)) / mp.sqrt(z) - 1) / z if z else mp.mpf(1) / 3, 0, mp.mpf("0.0295"), 6))
# exp(r) = 1 + r + r^2 E(r), r in [-ln2/2, ln2/2]
h = mp.log(2) / 2 + mp.mpf("1e-4")
show("E", *cheb_fit(lambda r: (mp.exp(r) - 1 - r) / r ** 2 if abs(r) > mp.mpf('1e-12') else mp.mpf(1) / 2 + r / 6, -h, h, 10))
