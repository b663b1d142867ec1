"""Coefficients of the table-driven device 2^y (openpbe_b200/csrc/pbe_math.cuh, exp2_neg_tab).

2^y = 2^k T[j] 2^f,  16 y = 16 k + j + 16 f,  j in 0..15,  |f| <= 1/32,  T[j] = 2^(j/16).
2^f ~ 1 + f s(f), s of degree `deg` (argv[1], default 5: total degree 6) by Chebyshev
interpolation of (2^f - 1)/f on |f| <= 1/32 (+1 % margin) in 60-digit arithmetic, rounded to
double.  Prints the coefficients, the table, and the max relative error of the double-rounded
polynomial evaluated exactly.
"""
import sys

import mpmath as mp

mp.mp.dps = 60
a = mp.mpf(1) / 32 * mp.mpf("1.01")
deg = int(sys.argv[1]) if len(sys.argv) > 1 else 5


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
print("coefficients c0.. (double):")
for k, c in enumerate(coef):
    print(f"  c{k:<2d} = {c!r}")
print("table T[j] = 2^(j/16):")
print(", ".join(repr(float(mp.mpf(2) ** (mp.mpf(j) / 16))) for j in range(16)))
worst = mp.mpf(0)
M = 4001
for i in range(M):
    f = mp.mpf(float((-mp.mpf(1) + 2 * mp.mpf(i) / (M - 1)) / 32))
    p = mp.mpf(0)
    for c in reversed(coef):
        p = p * f + mp.mpf(c)
    worst = max(worst, abs(p / mp.mpf(2) ** f - 1))
print("max rel error of the exact-arithmetic polynomial:", mp.nstr(worst, 5), "=",
      mp.nstr(worst / mp.mpf(2) ** -53, 5), "ulp(1)")
