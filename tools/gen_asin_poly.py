"""Generate the float64 polynomial used by usv_math.cuh::asin_k.

asin(x) = x + x^3 P(x^2) on |x| <= 0.5 (z = x^2 in [0, 0.25]); |x| > 0.5 goes through
asin(x) = pi/2 - 2 asin(sqrt((1 - |x|)/2)), which lands in the same domain.  P is the interpolant of
(asin(sqrt z)/sqrt z - 1)/z at Chebyshev nodes (near-minimax), computed with mpmath at 60 digits,
rounded to float64, and its error re-measured in float64 Horner/FMA order.

    python tools/gen_asin_poly.py            # prints the coefficients and the measured max error
"""
import mpmath as mp
import numpy as np

mp.mp.dps = 60
DEG = 12
ZMAX = mp.mpf("0.25")


def g(z):
    if z == 0:
        return mp.mpf(1) / 6
    s = mp.sqrt(z)
    return (mp.asin(s) / s - 1) / z


def main():
    n = DEG + 1
    nodes = [ZMAX / 2 * (1 + mp.cos(mp.pi * (2 * k + 1) / (2 * n))) for k in range(n)]
    A = mp.matrix(n, n)
    b = mp.matrix(n, 1)
    for i, z in enumerate(nodes):
        for j in range(n):
            A[i, j] = z ** j
        b[i] = g(z)
    c = mp.lu_solve(A, b)
    coef = [float(c[j]) for j in range(n)]
    print("// P(z) = c0 + c1 z + ... (tools/gen_asin_poly.py, degree %d)" % DEG)
    for j, v in enumerate(coef):
        print("    %s,  // c%d  %s" % (repr(v), j, float.hex(v)))
    # error of asin_k as the device evaluates it (Horner from the top; numpy float64 has no fma, so
    # this over-estimates the rounding slightly)
    xs = np.concatenate([np.linspace(0, 0.5, 200001), np.linspace(0.5, 1.0, 200001)])
    worst = 0.0
    for x in xs[:: 97]:
        a = abs(float(x))
        big = a > 0.5
        z = (1 - a) / 2 if big else a * a
        s = float(np.sqrt(z)) if big else a
        p = coef[-1]
        for v in coef[-2::-1]:
            p = p * z + v
        r = s + s * (z * p)
        if big:
            r = (np.pi / 2 - 2 * r) + 6.123233995736766e-17
        ref = mp.asin(mp.mpf(a))
        if ref != 0:
            ulp = np.spacing(float(ref))
            worst = max(worst, abs(float(mp.mpf(r) - ref)) / ulp)
    print("// max error over the sample: %.2f ulp" % worst)


if __name__ == "__main__":
    main()
