"""Independent NumPy restatement of the reference formulas.

TEST INFRASTRUCTURE ONLY.  Written separately from pbe_oracle.c (different language, vectorised
over cells, numpy.linalg.eigh instead of the hand-written QL) so that the two can be checked
against each other and against the reference's golden vectors.
"""
from __future__ import annotations

import numpy as np


# coalescenceKernels/CoulaloglouTavlarides.C:69-90
def coal_ct(c1, c2, gamma, sigma, xi1, xi2, eps, rho, nu, vcell=None, div_vcell=False):
    a = np.power(xi1, 1.0 / 3.0)
    b = np.power(xi2, 1.0 / 3.0)
    freq = (c1 * np.power(eps, 1.0 / 3.0) * np.power(a + b, 2)
            * np.power(np.power(xi1, 2.0 / 9.0) + np.power(xi2, 2.0 / 9.0), 0.5) / (1 + gamma))
    if div_vcell:
        freq = freq / vcell
    rate = np.exp(-c2 * rho * eps * (rho * nu) / (sigma ** 2 * (1 + gamma) ** 3)
                  * np.power(a * b / (a + b), 4.0))
    return freq * rate


# breakupKernels/CoulaloglouTavlarides.C:77-93
def break_ct(c1, c2, gamma, sigma, xi, rho_d, eps):
    return (c1 * np.power(eps, 1.0 / 3.0)
            * np.exp(-c2 * sigma * (1 + gamma) ** 2
                     / (rho_d * np.power(eps, 2.0 / 3.0) * np.power(xi, 5.0 / 9.0)))
            / ((1 + gamma) * np.power(xi, 2.0 / 9.0)))


def break_binary(xi):  # breakupKernels/binaryBreakup.C:51-53
    return xi * xi


def daughter_beta(xi1, xi2):  # uniformBinaryBreakup.C:55-60
    return (2.0 / xi2) * (xi2 - xi1 >= 0)


def daughter_moment(xi, k):  # uniformBinaryBreakup.C:62-69
    return 2 * np.power(xi, k) / (k + 1)


def _coal(spec, xa, xb, eps, rho, nu):
    kind, c = spec["coalescence"], spec.get("coal_c", ())
    if kind == "constantCoalescence":
        return np.full(np.broadcast(xa, xb, eps).shape, float(c[0]))
    if kind == "CoulaloglouTavlaridesC":
        return coal_ct(c[0], c[1], c[2], c[3], xa, xb, eps, rho, nu)
    return np.zeros(np.broadcast(xa, xb, eps).shape)


def _break(spec, xi, rho, eps):
    kind, c = spec["breakup"], spec.get("br_c", ())
    if kind == "binaryBreakup":
        return break_binary(xi) * np.ones_like(eps)
    if kind == "CoulaloglouTavlarides":
        return break_ct(c[0], c[1], c[2], c[3], xi, rho, eps)
    return np.zeros_like(eps)


def moc_source(spec, N, xi1, relax, n, eps, rho, nu, dt):
    """MOC.C:87-157 vectorised over cells; n is [N, ncells]."""
    xi = xi1 * (np.arange(N) + 1)
    g = np.stack([_break(spec, xi[i], rho, eps) for i in range(N)])
    S = np.empty_like(n)
    for i in range(N):
        f = np.zeros(n.shape[1])
        for j in range(N - i - 1):
            f -= _coal(spec, xi[i], xi[j], eps, rho, nu) * n[j]
        f *= n[i]
        for j in range(i):
            f += 0.5 * _coal(spec, xi[i - j - 1], xi[j], eps, rho, nu) * n[i - j - 1] * n[j]
        h = -g[i] * n[i]
        for j in range(i + 1, N):
            h += daughter_beta(xi[i], xi[j]) * g[j] * n[j] * xi1
        S[i] = np.maximum(f + h, -relax * n[i] / dt)
    return S


def wheeler(m):
    """MomentInversion.C:9-51 with numpy.linalg.eigh (the reference's own comment, :43)."""
    m = np.asarray(m, dtype=np.float64)
    N = m.size // 2
    sigma = np.zeros((N + 1, 2 * N))
    sigma[1, :] = m[: 2 * N]
    a = np.zeros(N)
    b = np.zeros(N)
    a[0] = m[1] / m[0]
    b[0] = m[0]
    for k in range(1, N):
        ncol = 2 * (N - k)
        sigma[k + 1, k:k + ncol] = (sigma[k, k + 1:k + 1 + ncol] - a[k - 1] * sigma[k, k:k + ncol]
                                    - b[k - 1] * sigma[k - 1, k:k + ncol])
        a[k] = -sigma[k, k] / sigma[k, k - 1] + sigma[k + 1, k + 1] / sigma[k + 1, k]
        b[k] = sigma[k + 1, k] / sigma[k, k - 1]
    bd = -np.sqrt(np.abs(b))
    J = np.diag(a)
    for i in range(N - 1):
        J[i, i + 1] = J[i + 1, i] = bd[i + 1]
    lam, V = np.linalg.eigh(J)
    return lam, m[0] * V[0, :] ** 2


def qmom_source_cell(spec, order, m, eps, rho, nu):
    """QMOM.C:226-264 for one cell; returns the 2*order moment sources."""
    x, w = wheeler(m)
    out = np.zeros(2 * order)
    e, r, v = np.float64(eps), np.float64(rho), np.float64(nu)
    for k in range(2 * order):
        cs = 0.0
        bs = 0.0
        for i in range(order):
            death = 0.0
            birth = 0.0
            for j in range(order):
                Sij = float(_coal(spec, x[i], x[j], e, r, v))
                death += w[j] * Sij
                birth += w[j] * np.power(x[i] + x[j], float(k)) * Sij
            cs += w[i] * (birth / 2.0 - np.power(x[i], float(k)) * death)
            bs += w[i] * float(_break(spec, x[i], r, e)) * (daughter_moment(x[i], k) - np.power(x[i], float(k)))
        out[k] = cs + bs
    return out


def transport_upwind_explicit(owner, neighbour, boundary_owner, vcell, psi, phi, dt, psi_b=None, S=None,
                              coupled=(0, 0), psi_halo=None):
    """Independent restatement of MULES::explicitSolve with the upwind flux (see pbe_oracle.h for the
    OpenFOAM functions it follows).  np.add.at accumulates unbuffered in index order = face order,
    so the result is bit-identical to the C face loop."""
    own, nei, bown = (np.asarray(a, dtype=np.int64) for a in (owner, neighbour, boundary_owner))
    psi = np.asarray(psi, dtype=np.float64)
    ni, nb = own.size, bown.size
    rDeltaT = 1.0 / dt
    out = np.empty_like(psi)
    c0, ncpl = coupled
    for k in range(psi.shape[0]):
        f = psi[k]
        lam = np.where(phi[:ni] >= 0.0, 1.0, 0.0)
        face = lam * (f[own] - f[nei]) + f[nei]
        flux = phi[:ni] * face
        # interleave += on the owner and -= on the neighbour in face order
        idx = np.empty(2 * ni, dtype=np.int64); idx[0::2] = own; idx[1::2] = nei
        val = np.empty(2 * ni); val[0::2] = flux; val[1::2] = -flux
        ivf = np.zeros(f.size)
        np.add.at(ivf, idx, val)
        bval = f[bown].copy() if psi_b is None else np.asarray(psi_b[k], dtype=np.float64).copy()
        if ncpl:
            w = np.where(phi[ni + c0:ni + c0 + ncpl] >= 0.0, 1.0, 0.0)
            bval[c0:c0 + ncpl] = w * f[bown[c0:c0 + ncpl]] + (1.0 - w) * psi_halo[k]
        np.add.at(ivf, bown, phi[ni:ni + nb] * bval)
        Su = 0.0 if S is None else S[k]
        out[k] = (f * rDeltaT + Su - ivf / vcell) / rDeltaT
    return out
