"""Fixed-pivot method of classes on an arbitrary (e.g. geometric) class grid -- CPU restatement.

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's cpu legs).

PARITY UNPINNED.  The reference's MOC only has the uniform grid xi_i = (i + 1) xi1
(src/PBESystem/MOC.C:56,79): coalescence of classes i and j lands exactly on class i + j + 1
(MOC.C:115-121).  The only hook for anything else is the comment at MOC.C:147-148 that names Kumar and
Ramkrishna (1996).  This file restates THEIR fixed-pivot technique in the reference's conventions, so
that on the uniform grid every coalescence term is the reference's own (checked in
tests/test_fixed_pivot.py against numpy_restatement.moc_source):

  * pivots x_0 < x_1 < ... < x_{N-1}, N_i = class number density (MOC.H:66);
  * coalescence (MOC.C:92-124): an unordered pair (i <= j) with v = x_i + x_j <= x_{N-1} (the
    reference drops pairs that leave the grid, MOC.C:107-109) dies at rate S_ij N_i N_j in both classes
    (a pair i == i once, with N_i^2) and is born (half of it for i == j: the 0.5 of MOC.C:118) on the
    two pivots that bracket v, x_t <= v < x_{t+1}, with weights eta = (x_{t+1} - v) / (x_{t+1} - x_t)
    and 1 - eta: number and volume of the new drop are both preserved;
  * breakup (MOC.C:126-157): class k dies at g_k N_k and its daughters, distributed with
    beta(v, x_k) = 2 / x_k on (0, x_k) (uniformBinaryBreakup.C:55-60), are assigned to the pivots through
    the same hat functions, n_{i,k} = int hat_i(v) beta(v, x_k) dv with x_{-1} = 0:
        k > i:  (2 / x_k) (x_{i+1} - x_{i-1}) / 2,      k == i:  (x_i - x_{i-1}) / x_i.
    Volume is conserved exactly; on the uniform grid the k > i weights are the reference's
    beta(x_i, x_k) deltaXi (MOC.C:146-154), the k == i term is what the reference's mean-value
    quadrature leaves out (it loses deltaXi of volume per breakup).
  * limiter max(S, -relax N_i / dt) as MOC.C:87-91.
"""
from __future__ import annotations

import numpy as np

from . import numpy_restatement as npo


SNAP = 8 * np.finfo(float).eps  # pbe_moc_create_pivots uses the same


def pair_list(x):
    """(i, j, t, eta) of every unordered pair that stays on the grid, i-major."""
    x = np.asarray(x, float)
    N = x.size
    out = []
    for i in range(N):
        for j in range(i, N):
            v = x[i] + x[j]
            vt = v * (1.0 + SNAP)  # a sum within a few ulp of a pivot is ON that pivot
            if v > x[N - 1] * (1.0 + SNAP):
                break
            t = min(int(np.searchsorted(x, vt, side="right")) - 1, N - 1)
            eta = 1.0 if t == N - 1 else min(1.0, max(0.0, (x[t + 1] - v) / (x[t + 1] - x[t])))
            out.append((i, j, t, eta))
    return out


def breakup_weights(x):
    """Hh_i = (x_{i+1} - x_{i-1}) / 2 (x_{-1} = 0; unused for the last class), Cself_i = (x_i - x_{i-1}) / x_i."""
    x = np.asarray(x, float)
    N = x.size
    xm = np.concatenate([[0.0], x[:-1]])
    xp = np.concatenate([x[1:], [x[-1]]])
    return 0.5 * (xp - xm), (x - xm) / x


def moc_source(spec, x, relax, n, eps, rho, nu, dt, limiter=True):
    """Class sources [N, ncells] on the pivots x; returns (S, scale) where scale is the sum of the
    magnitudes of every term that entered S_i (the tolerance unit of the device tests)."""
    x = np.asarray(x, float)
    N = x.size
    coal = np.zeros_like(n)
    scale = np.zeros_like(n)
    if spec["coalescence"] != "noCoalescence":
        for i, j, t, eta in pair_list(x):
            tt = npo._coal(spec, x[i], x[j], eps, rho, nu) * n[i] * n[j]
            if i == j:
                tt = 0.5 * tt
            coal[i] -= tt
            coal[j] -= tt
            coal[t] += eta * tt
            scale[i] += np.abs(tt)
            scale[j] += np.abs(tt)
            scale[t] += np.abs(eta * tt)
            if t + 1 < N:
                coal[t + 1] += (1.0 - eta) * tt
                scale[t + 1] += np.abs((1.0 - eta) * tt)
    br = np.zeros_like(n)
    if spec["breakup"] != "noBreakup":
        Hh, Cs = breakup_weights(x)
        gn = np.stack([npo._break(spec, x[k], rho, eps) * n[k] for k in range(N)])
        R = np.zeros(n.shape[1])
        for i in range(N - 1, -1, -1):
            br[i] = Hh[i] * R + (Cs[i] - 1.0) * gn[i]
            scale[i] += np.abs(Hh[i] * R) + np.abs(gn[i])
            R = R + (2.0 / x[i]) * gn[i]
    S = coal + br
    if limiter:
        S = np.maximum(S, -relax * n / dt)
    return S, scale


def geometric_pivots(N, x0, ratio):
    return x0 * ratio ** np.arange(N)
