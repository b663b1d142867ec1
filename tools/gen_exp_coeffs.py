"""Derive the degree-11 polynomial used by the device exp (openpbe_b200/csrc/pbe_math.cuh).

exp(r) ~ 1 + r + r^2 q(r) on |r| <= ln2/2 (+1% margin), q of degree 9 obtained by Chebyshev
interpolation of (exp(r) - 1 - r)/r^2 in 60-digit arithmetic (near-minimax), rounded to
double.  Prints the coefficients and the max relative error of the double-rounded polynomial
evaluated exactly, so the header can quote it.
"""
import mpmath as mp

mp.mp.dps = 60
a = mp.log(2) / 2 * mp.mpf("1.01")
deg_q = 9


def g(r):
    if abs(r) < mp.mpf("1e-25"):
        return mp.mpf(1) / 2 + r / 6
    return (mp.e ** r - 1 - r) / (r * r)


nodes = [a * mp.cos(mp.pi * (2 * k + 1) / (2 * (deg_q + 1))) for k in range(deg_q + 1)]
A = mp.matrix(deg_q + 1, deg_q + 1)
b = mp.matrix(deg_q + 1, 1)
for i, x in enumerate(nodes):
    for j in range(deg_q + 1):
        A[i, j] = x ** j
    b[i] = g(x)
q = mp.lu_solve(A, b)
coef = [1.0, 1.0] + [float(q[j]) for j in range(deg_q + 1)]
print("coefficients c0..c11 (double):")
for k, c in enumerate(coef):
    print(f"  c{k:<2d} = {c!r}  ({float(c).hex()})")
worst = mp.mpf(0)
M = 4001
for i in range(M):
    r = -mp.log(2) / 2 + mp.log(2) * i / (M - 1)
    r = mp.mpf(float(r))
    p = mp.mpf(0)
    for c in reversed(coef):
        p = p * r + mp.mpf(c)
    err = abs(p / mp.e ** r - 1)
    worst = max(worst, err)
print("max rel error of the exact-arithmetic polynomial:", mp.nstr(worst, 5),
      "=", mp.nstr(worst / mp.mpf(2) ** -53, 5), "ulp(1)")
