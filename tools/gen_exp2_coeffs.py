"""Derive the polynomial of the device 2^f (openpbe_b200/csrc/pbe_math.cuh, exp2_frac).

2^f ~ 1 + f s(f) on |f| <= 1/2 (+1% margin), s of degree 9 (argv[1]) by Chebyshev interpolation of
(2^f - 1)/f in 60-digit arithmetic (near-minimax), rounded to double.  Prints the coefficients
and the max relative error of the double-rounded polynomial evaluated exactly.
"""
import sys

import mpmath as mp

mp.mp.dps = 60
a = mp.mpf("0.505")
deg = int(sys.argv[1]) if len(sys.argv) > 1 else 9  # degree of s; the device uses 9 (5.7 ulp)


def g(f):
    if abs(f) < mp.mpf("1e-25"):
        return mp.log(2) + f * mp.log(2) ** 2 / 2
    return (mp.mpf(2) ** f - 1) / f


nodes = [a * mp.cos(mp.pi * (2 * k + 1) / (2 * (deg + 1))) for k in range(deg + 1)]
A = mp.matrix(deg + 1, deg + 1)
b = mp.matrix(deg + 1, 1)
for i, x in enumerate(nodes):
    for j in range(deg + 1):
        A[i, j] = x ** j
    b[i] = g(x)
q = mp.lu_solve(A, b)
coef = [1.0] + [float(q[j]) for j in range(deg + 1)]
print("coefficients c0..c11 (double):")
for k, c in enumerate(coef):
    print(f"  c{k:<2d} = {c!r}")
worst = mp.mpf(0)
M = 4001
for i in range(M):
    f = mp.mpf(float(-mp.mpf(1) / 2 + mp.mpf(i) / (M - 1)))
    p = mp.mpf(0)
    for c in reversed(coef):
        p = p * f + mp.mpf(c)
    worst = max(worst, abs(p / mp.mpf(2) ** f - 1))
print("max rel error of the exact-arithmetic polynomial:", mp.nstr(worst, 5), "=",
      mp.nstr(worst / mp.mpf(2) ** -53, 5), "ulp(1)")
